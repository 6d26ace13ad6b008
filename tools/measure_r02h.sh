cd $GRAFT_REPO_ROOT
O=gpurun_out
for spc in 2 3 6 8; do NNBVH_OUT=/tmp/libnnbvh_spc$spc.so NNBVH_EXTRA_FLAGS=-DNNBVH_SPC=$spc sh nn_bvh_b200/csrc/build.sh > /dev/null 2>&1 & done; wait
python tools/sweep.py bathroom 8388608 NNBVH_KERNEL=3 > $O/r02h_sweep.txt 2>&1
for spc in 2 3 6 8; do echo "SPC=$spc" >> $O/r02h_sweep.txt; NNBVH_LIB=/tmp/libnnbvh_spc$spc.so python tools/sweep.py bathroom 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02h_sweep.txt; done
python tools/sweep.py crown 8388608 NNBVH_KERNEL=3 >> $O/r02h_sweep.txt 2>&1
for spc in 2 8; do echo "SPC=$spc" >> $O/r02h_sweep.txt; NNBVH_LIB=/tmp/libnnbvh_spc$spc.so python tools/sweep.py crown 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02h_sweep.txt; done
cat $O/r02h_sweep.txt
