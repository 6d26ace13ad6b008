# final: the GPU suite on two GPUs (the peer-window test needs two), smoke
cd $GRAFT_REPO_ROOT
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > $O/r02zt_tests_n2.log; cat $O/r02zt_tests_n2.log
python __graft_entry__.py smoke 2>&1 | tail -1
