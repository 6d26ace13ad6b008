cd $GRAFT_REPO_ROOT
O=gpurun_out
: > $O/r02zl_sweep.txt
for sc in bathroom crown; do
echo "pred-pop (default) $sc" >> $O/r02zl_sweep.txt; NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_prev.so python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zl_sweep.txt
echo "branch pop $sc" >> $O/r02zl_sweep.txt; NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_alt.so python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zl_sweep.txt
done
cat $O/r02zl_sweep.txt
