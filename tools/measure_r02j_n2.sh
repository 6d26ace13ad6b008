cd $GRAFT_REPO_ROOT
O=gpurun_out
python -m pytest tests/test_gpu_peer_window.py -m gpu -x -q 2>&1 | tail -3 > $O/r02j_peer_test.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > $O/r02j_bench_n2.json 2> $O/r02j_bench_n2.err; echo "rc=$?" >> $O/r02j_bench_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/host_link_probe.py > $O/r02j_host_link_n2.json 2> $O/r02j_host_link_n2.err
cat $O/r02j_peer_test.log; tail -3 $O/r02j_bench_n2.err; cat $O/r02j_host_link_n2.json; tail -2 $O/r02j_host_link_n2.err
