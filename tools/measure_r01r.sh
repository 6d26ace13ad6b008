set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out
timeout 60 python -m pytest tests/test_wavefront.py -x -q -m gpu 2>&1 | tail -4
timeout 110 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 12 -c 4 -f -o $O/r01r_step python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r01r_ncu_full.out 2>&1
ls -la $O | tail -5
