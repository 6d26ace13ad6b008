cd $GRAFT_REPO_ROOT
O=gpurun_out
for pf in 4096 16384 65536; do NNBVH_OUT=/tmp/libnnbvh_pf$pf.so NNBVH_EXTRA_FLAGS=-DNNBVH_PREFETCH_AHEAD=$pf sh nn_bvh_b200/csrc/build.sh > /dev/null 2>&1 & done; wait
: > $O/r02o_sweep.txt
python tools/sweep.py bathroom 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02o_sweep.txt
for pf in 4096 16384 65536; do echo "PREFETCH_AHEAD=$pf" >> $O/r02o_sweep.txt; NNBVH_LIB=/tmp/libnnbvh_pf$pf.so python tools/sweep.py bathroom 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02o_sweep.txt; done
python tools/sweep.py crown 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02o_sweep.txt
for pf in 16384; do echo "PREFETCH_AHEAD=$pf crown" >> $O/r02o_sweep.txt; NNBVH_LIB=/tmp/libnnbvh_pf$pf.so python tools/sweep.py crown 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02o_sweep.txt; done
cat $O/r02o_sweep.txt
