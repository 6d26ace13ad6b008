cd $GRAFT_REPO_ROOT
O=gpurun_out
: > $O/r02zs_sweep.txt
for sc in bathroom crown; do
echo "$sc" >> $O/r02zs_sweep.txt; python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 NNBVH_KERNEL=3,NNBVH_L2_PERSIST_MB=24,NNBVH_VERBOSE=1 NNBVH_KERNEL=3,NNBVH_L2_PERSIST_MB=40 NNBVH_KERNEL=3,NNBVH_L2_PERSIST_MB=64 NNBVH_KERNEL=3,NNBVH_L2_PERSIST_MB=96 2> $O/r02zs_$sc.err | tail -5 >> $O/r02zs_sweep.txt
grep "nnbvh:" $O/r02zs_$sc.err | head -1 >> $O/r02zs_sweep.txt
done
cat $O/r02zs_sweep.txt
