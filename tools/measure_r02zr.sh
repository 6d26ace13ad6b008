cd $GRAFT_REPO_ROOT
O=gpurun_out
: > $O/r02zr_sweep.txt
for sc in bathroom crown; do
echo "default $sc" >> $O/r02zr_sweep.txt; python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zr_sweep.txt
echo "tri tex $sc" >> $O/r02zr_sweep.txt; NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_alt.so python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zr_sweep.txt
done
cat $O/r02zr_sweep.txt
NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_alt.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -x -q 2>&1 | tail -3
