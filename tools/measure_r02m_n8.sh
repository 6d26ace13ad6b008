# round 2, 8 GPUs: host-link probe, the full bench line, config-5 exchange variants, config-5 at N=1/4 on the same box
cd $GRAFT_REPO_ROOT
O=gpurun_out
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$T --nproc-per-node 8 --master-port 29531 tools/host_link_probe.py > $O/r02m_host_link_n8.json 2> $O/r02m.err
$T --nproc-per-node 8 --master-port 29532 bench.py --gpus 8 --steps 10 --warmup 3 > $O/r02m_bench_n8.json 2>> $O/r02m.err; echo "bench rc=$?" >> $O/r02m.err
run() { $T --nproc-per-node $1 --master-port $2 bench.py --gpus $1 --workload config5 --steps 3 --c5-steps 3 $3 2>> $O/r02m.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); c=d['configs']['synthetic10M']
print('N=$1 $3', round(c['value'],1), round(c['ms_per_step'],2), 'kernel_only', round(c['kernel_only_ms'],2), c['checksum']['ok'], c['checksum']['gathered_records_match'])
" >> $O/r02m_c5_variants.txt; }
: > $O/r02m_c5_variants.txt
run 8 29541 "--c5-lanes 1 --c5-chunks 1"
run 8 29542 "--c5-lanes 2 --c5-chunks 0.5,0.5"
run 8 29543 "--c5-lanes 2 --c5-chunks 0.4,0.3,0.2,0.1"
run 8 29544 "--c5-lanes 2 --c5-chunks 0.4,0.3,0.3 --c5-rank0-share 0.8"
run 8 29545 "--c5-lanes 1 --c5-chunks 1 --c5-rank0-share 0.7"
run 8 29546 "--c5-lanes 2 --c5-chunks 0.5,0.5 --c5-rank0-share 0.6"
run 4 29547 ""
python bench.py --workload config5 --steps 3 --c5-steps 3 2>> $O/r02m.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); c=d['configs']['synthetic10M']
print('N=1', round(c['value'],1), round(c['ms_per_step'],2), 'kernel_only', round(c['kernel_only_ms'],2), c['checksum']['ok'])
" >> $O/r02m_c5_variants.txt
cat $O/r02m_c5_variants.txt; cat $O/r02m_host_link_n8.json; grep -v "OMP_NUM\|\*\*\*\*" $O/r02m.err | tail -5
