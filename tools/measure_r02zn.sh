# round 2, closing single-GPU evidence after the StepKernel rework (predicated pop, no lane-carried hit record):
# bench (all workloads) + reference arm + launch list + ncu --set full of three workloads + side benches
cd $GRAFT_REPO_ROOT
O=gpurun_out
python bench.py --steps 20 --warmup 5 > $O/r02zn_bench.json 2> $O/r02zn_bench.err; echo "rc=$?" >> $O/r02zn_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r02zn_bench_reference.json 2>> $O/r02zn_bench.err
python bench.py --workload config2 --scene crown --steps 20 --warmup 5 --no-cpu-baseline > $O/r02zn_bench_crown.json 2>> $O/r02zn_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/r02zn_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r02zn_ncu_launches.out 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 12 -c 4 -f -o $O/r02zn_step_bathroom python bench.py --workload config2 --steps 2 --warmup 3 --no-cpu-baseline > $O/r02zn_ncu_bathroom.out 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 12 -c 4 -f -o $O/r02zn_step_crown python bench.py --workload config2 --scene crown --steps 2 --warmup 3 --no-cpu-baseline > $O/r02zn_ncu_crown.out 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 4 -c 2 -f -o $O/r02zn_step_config5 python bench.py --workload config5 --steps 1 --c5-steps 1 > $O/r02zn_ncu_config5.out 2>&1
python __graft_entry__.py smoke > $O/r02zn_smoke.log 2>&1
python tools/instances_bench.py > $O/r02zn_instances_bench.txt 2>&1
python tools/alpha_bench.py > $O/r02zn_alpha_bench.txt 2>&1
python tools/config3.py > $O/r02zn_config3_crown.jsonl 2> $O/r02zn_config3.err
ls -la $O | grep r02zn; tail -2 $O/r02zn_bench.err; tail -1 $O/r02zn_smoke.log; cat $O/r02zn_bench.json
