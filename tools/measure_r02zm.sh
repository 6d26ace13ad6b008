cd $GRAFT_REPO_ROOT
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $O/r02zm_tests.log; cat $O/r02zm_tests.log
: > $O/r02zm_sweep.txt
for sc in bathroom crown; do
echo "default (SPC 4) $sc" >> $O/r02zm_sweep.txt; python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 NNBVH_KERNEL=3,NNBVH_WIDE_MINB=6 2>&1 | tail -2 >> $O/r02zm_sweep.txt
for v in 3 6; do echo "SPC $v $sc" >> $O/r02zm_sweep.txt; NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_spc$v.so python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02zm_sweep.txt; done
done
cat $O/r02zm_sweep.txt
