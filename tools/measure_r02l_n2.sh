cd $GRAFT_REPO_ROOT
O=gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --workload config5 --steps 3 --c5-steps 3 $2 2>> $O/r02l.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); c=d['configs']['synthetic10M']
print('$2', round(c['value'],1), round(c['ms_per_step'],2), 'kernel_only', round(c['kernel_only_ms'],2), c['checksum']['ok'], c['checksum']['gathered_records_match'])
" >> $O/r02l_c5_variants.txt; }
: > $O/r02l_c5_variants.txt
run 29521 "--c5-lanes 1 --c5-chunks 1 --c5-copy-streams 1"
run 29522 "--c5-lanes 1 --c5-chunks 1 --c5-copy-streams 2"
run 29523 "--c5-lanes 1 --c5-chunks 1 --c5-copy-streams 4"
run 29524 "--c5-lanes 2 --c5-chunks 0.5,0.5 --c5-copy-streams 4"
run 29525 "--c5-lanes 2 --c5-chunks 0.7,0.3 --c5-copy-streams 4"
python bench.py --workload config5 --steps 3 --c5-steps 3 2>> $O/r02l.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); c=d['configs']['synthetic10M']
print('N=1', round(c['value'],1), round(c['ms_per_step'],2), 'kernel_only', round(c['kernel_only_ms'],2), c['checksum']['ok'], 'roofline', c['roofline']['bound'], round(c['roofline']['frac'],3))
" >> $O/r02l_c5_variants.txt
cat $O/r02l_c5_variants.txt; tail -3 $O/r02l.err
