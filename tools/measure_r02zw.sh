cd $GRAFT_REPO_ROOT
O=gpurun_out
: > $O/r02zw_sweep.txt
for sc in bathroom crown; do
echo "default $sc" >> $O/r02zw_sweep.txt; python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zw_sweep.txt
echo "refill window $sc" >> $O/r02zw_sweep.txt; NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_win.so python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 NNBVH_KERNEL=3,NNBVH_REFILL_BELOW=16 NNBVH_KERNEL=3,NNBVH_REFILL_BELOW=20 NNBVH_KERNEL=3,NNBVH_REFILL_BELOW=24 NNBVH_KERNEL=3,NNBVH_REFILL_BELOW=28 2>&1 | tail -5 >> $O/r02zw_sweep.txt
done
cat $O/r02zw_sweep.txt
NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_win.so timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -x -q 2>&1 | tail -3
