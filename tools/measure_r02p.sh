cd $GRAFT_REPO_ROOT
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > $O/r02p_gputest.log
python tools/sweep.py bathroom 8388608 NNBVH_KERNEL=3 > $O/r02p_sweep.txt 2>&1
python tools/sweep.py crown 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02p_sweep.txt
python tools/sweep.py killeroos 8388608 NNBVH_KERNEL=3 2>&1 | tail -1 >> $O/r02p_sweep.txt
python bench.py --workload config2 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02p_bench.json 2> $O/r02p_bench.err
python bench.py --workload config2 --scene crown --steps 10 --warmup 3 --no-cpu-baseline > $O/r02p_bench_crown.json 2>> $O/r02p_bench.err
cat $O/r02p_gputest.log $O/r02p_sweep.txt; python -c "
import json
for f in ('r02p_bench.json','r02p_bench_crown.json'):
    d=json.load(open('$O/'+f)); print(f, round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],3), round(d['roofline']['kernel_ms'],3), d['roofline']['bound'], round(d['roofline']['frac'],3))
"
