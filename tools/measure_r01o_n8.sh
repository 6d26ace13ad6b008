# 8-GPU box: BASELINE.json configs[4] at 1/4/8 GPUs with the hit records returned to rank 0, and bench.py at N=8
cd $GRAFT_REPO_ROOT
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
python tools/config5.py --gather none > $O/r01o_config5_n1.txt 2>&1
$TR --nproc-per-node 4 --master-port 29514 tools/config5.py --gather peer > $O/r01o_config5_n4.txt 2>&1
$TR --nproc-per-node 8 --master-port 29518 tools/config5.py --gather peer > $O/r01o_config5_n8.txt 2>&1
$TR --nproc-per-node 8 --master-port 29519 tools/config5.py --gather nccl > $O/r01o_config5_n8_nccl.txt 2>&1
$TR --nproc-per-node 8 --master-port 29520 bench.py --gpus 8 --steps 30 --warmup 3 > $O/r01o_bench_n8.json 2> $O/r01o_bench_n8.err
tail -n 2 $O/r01o_config5_n1.txt $O/r01o_config5_n4.txt $O/r01o_config5_n8.txt $O/r01o_config5_n8_nccl.txt
tail -c 600 $O/r01o_bench_n8.json
