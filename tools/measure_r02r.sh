cd $GRAFT_REPO_ROOT
O=gpurun_out
for v in 1 2; do NNBVH_OUT=/tmp/libnnbvh_pt$v.so NNBVH_EXTRA_FLAGS=-DNNBVH_PREFETCH_TRI=$v sh nn_bvh_b200/csrc/build.sh > /dev/null 2>&1 & done; wait
: > $O/r02r_sweep.txt
for sc in bathroom crown; do
python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02r_sweep.txt
for v in 1 2; do echo "PREFETCH_TRI=$v $sc" >> $O/r02r_sweep.txt; NNBVH_LIB=/tmp/libnnbvh_pt$v.so python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02r_sweep.txt; done
done
cat $O/r02r_sweep.txt
