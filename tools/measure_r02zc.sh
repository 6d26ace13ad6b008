cd $GRAFT_REPO_ROOT
O=gpurun_out
: > $O/r02zc_sweep.txt
CF=""
for q in 6 8 10 12; do for r in 8 12 16 20; do CF="$CF NNBVH_KERNEL=3,NNBVH_LEAF_QUORUM=$q,NNBVH_REFILL_BELOW=$r"; done; done
for sc in bathroom crown; do
echo "$sc" >> $O/r02zc_sweep.txt; python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 $CF 2>&1 | tail -17 >> $O/r02zc_sweep.txt
done
cat $O/r02zc_sweep.txt
