# round 2, step c: new bench line (config 2 + config 5), reference arm, launch list, ncu --set full of both workloads
cd $GRAFT_REPO_ROOT
O=gpurun_out
python bench.py --steps 10 --warmup 3 > $O/r02c_bench.json 2> $O/r02c_bench.err; echo "rc=$?" >> $O/r02c_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/r02c_bench_reference.json 2>> $O/r02c_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/r02c_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r02c_ncu_launches.out 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 12 -c 4 -f -o $O/r02c_step_bathroom python bench.py --workload config2 --steps 2 --warmup 3 --no-cpu-baseline > $O/r02c_ncu_full_bathroom.out 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 4 -c 2 -f -o $O/r02c_step_config5 python bench.py --workload config5 --steps 1 --c5-steps 1 > $O/r02c_ncu_full_config5.out 2>&1
ls -la $O | tail -8; tail -3 $O/r02c_bench.err
