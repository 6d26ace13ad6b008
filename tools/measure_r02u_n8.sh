# round 2, final 8-GPU session: config-5 chunk schedules, then the full bench line at N=8 (and N=1 config 5 on the same box)
cd $GRAFT_REPO_ROOT
O=gpurun_out
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { $T --nproc-per-node $1 --master-port $2 bench.py --gpus $1 --workload config5 --steps 3 --c5-steps 3 $3 2>> $O/r02u.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); c=d['configs']['synthetic10M']
print('N=$1 $3', round(c['value'],1), round(c['ms_per_step'],2), 'kernel_only', round(c['kernel_only_ms'],2), c['checksum']['ok'], c['checksum']['gathered_records_match'])
" >> $O/r02u_c5_variants.txt; }
: > $O/r02u_c5_variants.txt
run 8 29541 "--c5-chunks 0.5,0.5"
run 8 29542 "--c5-chunks 0.45,0.45,0.1"
run 8 29543 "--c5-chunks 0.5,0.4,0.1"
run 8 29544 "--c5-chunks 0.6,0.3,0.1"
run 8 29545 "--c5-chunks 0.4,0.4,0.2"
run 8 29546 "--c5-chunks 0.55,0.45"
python bench.py --workload config5 --steps 3 --c5-steps 3 2>> $O/r02u.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); c=d['configs']['synthetic10M']
print('N=1', round(c['value'],1), round(c['ms_per_step'],2), 'kernel_only', round(c['kernel_only_ms'],2), c['checksum']['ok'])
" >> $O/r02u_c5_variants.txt
BEST=$(sort -k4,4 -n -r $O/r02u_c5_variants.txt | grep "N=8" | head -1 | awk '{print $3}')
echo "best chunks: $BEST" >> $O/r02u_c5_variants.txt
$T --nproc-per-node 8 --master-port 29532 bench.py --gpus 8 --steps 10 --warmup 3 --c5-chunks $BEST > $O/r02u_bench_n8.json 2>> $O/r02u.err; echo "bench rc=$?" >> $O/r02u.err
cat $O/r02u_c5_variants.txt; grep -v "OMP_NUM\|\*\*\*\*\|NCCL version\|^$" $O/r02u.err | tail -5
