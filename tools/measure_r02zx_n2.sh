# closing 2-GPU line on the final build
cd $GRAFT_REPO_ROOT
O=gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > $O/r02zx_bench_n2.json 2> $O/r02zx.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02zx_bench_n2.json").read().strip().split("\n")[-1]); c=d["configs"]["synthetic10M"]
print("N=2 value", round(d["value"]), "ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"]), "| config5", round(c["value"]), round(c["ms_per_step"],2), "kernel_only", round(c["kernel_only_ms"],2), c["checksum"]["ok"], c["checksum"]["gathered_records_match"])
PY
