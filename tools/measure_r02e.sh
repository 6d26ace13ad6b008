# round 2, step e: any-hit stack specialisation + select hit update; parity subset, knob sweep
cd $GRAFT_REPO_ROOT
O=gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py tests/test_spheres.py -m gpu -x -q 2>&1 | tail -3 > $O/r02e_gputest.log
python tools/sweep.py bathroom 8388608 NNBVH_KERNEL=3 NNBVH_KERNEL=3,NNBVH_LEAF_QUORUM=8 NNBVH_KERNEL=3,NNBVH_LEAF_QUORUM=10 NNBVH_KERNEL=3,NNBVH_LEAF_QUORUM=14 NNBVH_KERNEL=3,NNBVH_LEAF_QUORUM=16 NNBVH_KERNEL=3,NNBVH_REFILL_BELOW=16 NNBVH_KERNEL=3,NNBVH_REFILL_BELOW=24 NNBVH_KERNEL=3,NNBVH_REFILL_BELOW=28 NNBVH_KERNEL=3,NNBVH_LEAF_QUORUM=14,NNBVH_REFILL_BELOW=24 NNBVH_KERNEL=3,NNBVH_LEAF_QUORUM=10,NNBVH_REFILL_BELOW=16 NNBVH_KERNEL=3,NNBVH_WIDE_MINB=4 NNBVH_KERNEL=3,NNBVH_WIDE_MINB=6 > $O/r02e_sweep.txt 2>&1
cat $O/r02e_gputest.log; cat $O/r02e_sweep.txt
