cd $GRAFT_REPO_ROOT
O=gpurun_out
: > $O/r02zk_sweep.txt
for sc in bathroom crown; do
echo "new $sc" >> $O/r02zk_sweep.txt; python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zk_sweep.txt
echo "prev $sc" >> $O/r02zk_sweep.txt; NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_prev.so python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zk_sweep.txt
done
cat $O/r02zk_sweep.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $O/r02zk_tests.log; cat $O/r02zk_tests.log
