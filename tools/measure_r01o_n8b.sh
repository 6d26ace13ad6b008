cd $GRAFT_REPO_ROOT
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29521 tools/config5.py --gather direct > $O/r01o_config5_n8_direct.txt 2>&1
$TR --nproc-per-node 8 --master-port 29522 tools/config5.py --gather peer --chunks 3 > $O/r01o_config5_n8_peer3.txt 2>&1
tail -n 1 $O/r01o_config5_n8_direct.txt $O/r01o_config5_n8_peer3.txt
