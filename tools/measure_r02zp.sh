# final check of the committed build: the GPU suite as the driver runs it, smoke, one default bench line
cd $GRAFT_REPO_ROOT
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > $O/r02zp_tests.log; cat $O/r02zp_tests.log
python __graft_entry__.py smoke 2>&1 | tail -1
python bench.py > $O/r02zp_bench.json 2> $O/r02zp_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02zp_bench.json").read().strip().split("\n")[-1]); c=d["configs"]["synthetic10M"]
print("value", round(d["value"]), "ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"]), "launches", d["gpu_launches"], "clocks", d["clocks"], "roofline", d["roofline"]["bound"], round(d["roofline"]["frac"],3), "cpu", round(d["cpu_baseline"]["value"],1), "| config5", round(c["value"]), c["checksum"]["ok"])
PY
