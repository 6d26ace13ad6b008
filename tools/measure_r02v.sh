cd $GRAFT_REPO_ROOT
O=gpurun_out
NNBVH_OUT=/tmp/libnnbvh_ao0.so NNBVH_EXTRA_FLAGS=-DNNBVH_ANY_ORDER=0 sh nn_bvh_b200/csrc/build.sh > /dev/null 2>&1
: > $O/r02v_sweep.txt
for sc in bathroom crown; do
python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02v_sweep.txt
echo "ANY_ORDER=0 $sc" >> $O/r02v_sweep.txt; NNBVH_LIB=/tmp/libnnbvh_ao0.so python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02v_sweep.txt
done
cat $O/r02v_sweep.txt
