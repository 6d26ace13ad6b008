# round 2, closing 8-GPU line on the reworked StepKernel: the full bench at N=8
cd $GRAFT_REPO_ROOT
O=gpurun_out
timeout 380 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8 --master-port 29532 bench.py --gpus 8 --steps 10 --warmup 3 > $O/r02zi_bench_n8.json 2> $O/r02zi.err; echo "bench rc=$?" >> $O/r02zi.err
grep -v "OMP_NUM\|\*\*\*\*\|NCCL version\|^$" $O/r02zi.err | tail -3
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02zi_bench_n8.json").read().strip().split("\n")[-1]); c=d["configs"]["synthetic10M"]
print("N=8 value", round(d["value"]), "ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"]), "| config5", round(c["value"]), round(c["ms_per_step"],2), "kernel_only", round(c["kernel_only_ms"],2), c["checksum"]["ok"], c["checksum"]["gathered_records_match"])
PY
