cd $GRAFT_REPO_ROOT
O=gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 12 -c 4 -f -o $O/r02zd_step_bathroom python bench.py --workload config2 --steps 2 --warmup 3 --no-cpu-baseline > $O/r02zd_ncu_bathroom.out 2>&1
ls -la $O | grep r02zd; tail -2 $O/r02zd_ncu_bathroom.out
