set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out
python bench.py > $O/r01n_bench.json 2> $O/r01n_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r01n_bench_reference.json 2>> $O/r01n_bench.err
python bench.py --scene crown --steps 10 > $O/r01n_bench_crown.json 2>> $O/r01n_bench.err
python tools/scenes_table.py killeroos crown > $O/r01n_scenes.txt 2>&1
python tools/scenes_table.py killeroos crown --spheres > $O/r01n_scenes_spheres.txt 2>&1
python tools/nn_tree_bench.py > $O/r01n_nn_tree_bench.jsonl 2> $O/r01n_nn_tree_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r01n_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r01n_ncu_launches.out 2>&1
ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 12 -c 4 -f -o $O/r01n_step python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r01n_ncu_full.out 2>&1
ncu --set full --clock-control none --import-source on -k regex:StepKernel --launch-skip 4 -c 2 -f -o $O/r01n_step_crown python tools/prof.py crown 8388608 bounce 3 > $O/r01n_ncu_crown.out 2>&1
ls -la $O
