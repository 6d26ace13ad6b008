cd $GRAFT_REPO_ROOT
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/r01p_gpu_tests.log 2>&1; echo "rc=$?" >> $O/r01p_gpu_tests.log
python __graft_entry__.py smoke > $O/r01p_smoke.log 2>&1; echo "rc=$?" >> $O/r01p_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > $O/r01p_bench_reference.json 2> $O/r01p_bench.err
python bench.py > $O/r01p_bench.json 2>> $O/r01p_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r01p_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r01p_ncu_launches.out 2>&1
tail -n 3 $O/r01p_gpu_tests.log $O/r01p_smoke.log; tail -n 2 $O/r01p_bench.err
