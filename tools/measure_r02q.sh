# round 2, final single-GPU evidence: bench (all workloads) + reference arm + launch list + ncu --set full of three workloads
cd $GRAFT_REPO_ROOT
O=gpurun_out
python bench.py --steps 20 --warmup 5 > $O/r02q_bench.json 2> $O/r02q_bench.err; echo "rc=$?" >> $O/r02q_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r02q_bench_reference.json 2>> $O/r02q_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/r02q_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r02q_ncu_launches.out 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 12 -c 4 -f -o $O/r02q_step_bathroom python bench.py --workload config2 --steps 2 --warmup 3 --no-cpu-baseline > $O/r02q_ncu_bathroom.out 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 12 -c 4 -f -o $O/r02q_step_crown python bench.py --workload config2 --scene crown --steps 2 --warmup 3 --no-cpu-baseline > $O/r02q_ncu_crown.out 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 4 -c 2 -f -o $O/r02q_step_config5 python bench.py --workload config5 --steps 1 --c5-steps 1 > $O/r02q_ncu_config5.out 2>&1
python __graft_entry__.py smoke > $O/r02q_smoke.log 2>&1
ls -la $O | grep r02q; tail -2 $O/r02q_bench.err; tail -1 $O/r02q_smoke.log
