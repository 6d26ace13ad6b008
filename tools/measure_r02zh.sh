cd $GRAFT_REPO_ROOT
O=gpurun_out
for rep in 1 2; do
python bench.py --workload config2 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02zh_new$rep.json 2>/dev/null
NNBVH_LIB=$GRAFT_REPO_ROOT/nn_bvh_b200/libnnbvh_prev.so python bench.py --workload config2 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02zh_prev$rep.json 2>/dev/null
done
python - <<'PY'
import json
for t in ("new1","prev1","new2","prev2"):
    d=json.loads(open(f"gpurun_out/r02zh_{t}.json").read().strip().split("\n")[-1]); e=d["e2e"]
    print(t, round(d["value"]), round(e["value"]), round(e["one_lane_value"]), round(e["aos_value"]))
PY
