cd $GRAFT_REPO_ROOT
O=gpurun_out
: > $O/r02zu_sweep.txt
for sc in bathroom crown; do
echo "default $sc" >> $O/r02zu_sweep.txt; python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zu_sweep.txt
for v in 1 2; do echo "L1 no-allocate $v $sc" >> $O/r02zu_sweep.txt; NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_na$v.so python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zu_sweep.txt; done
done
cat $O/r02zu_sweep.txt
