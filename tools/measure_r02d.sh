# round 2, step d: branch-free node-step tail + permute; parity, bench, ncu counters of the two hot kernels
cd $GRAFT_REPO_ROOT
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $O/r02d_gputest.log
python bench.py --workload config2 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02d_bench.json 2> $O/r02d_bench.err; echo "rc=$?" >> $O/r02d_bench.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 12 -c 4 -f -o $O/r02d_step_bathroom python bench.py --workload config2 --steps 2 --warmup 3 --no-cpu-baseline > $O/r02d_ncu_full_bathroom.out 2>&1
python tools/sweep.py bathroom 8388608 NNBVH_KERNEL=3 > $O/r02d_sweep.txt 2>&1
cat $O/r02d_gputest.log; tail -2 $O/r02d_bench.err; cat $O/r02d_sweep.txt | tail -3
