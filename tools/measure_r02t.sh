cd $GRAFT_REPO_ROOT
O=gpurun_out
python -m pytest tests/test_general_scene.py tests/test_instances.py tests/test_alpha_gpu.py -m gpu -x -q 2>&1 | tail -4 > $O/r02t_tests.log
python tools/instances_bench.py > $O/r02t_instances_bench.txt 2>&1
cat $O/r02t_tests.log $O/r02t_instances_bench.txt
