cd $GRAFT_REPO_ROOT
O=gpurun_out
: > $O/r02zo_sweep.txt
CF=""
for b in 192 224 128 96 64; do CF="$CF NNBVH_KERNEL=3,NNBVH_STEP_BLOCK=$b"; done
for sc in bathroom crown; do
echo "$sc" >> $O/r02zo_sweep.txt; python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 $CF 2>&1 | tail -6 >> $O/r02zo_sweep.txt
done
cat $O/r02zo_sweep.txt
