cd $GRAFT_REPO_ROOT
O=gpurun_out
for v in "" "NNBVH_GRID_RAYS_PER_LANE=4" "NNBVH_GRID_RAYS_PER_LANE=8" "NNBVH_GRID_RAYS_PER_LANE=16" "NNBVH_BIN_MIN_RAYS=262144" "NNBVH_BIN_MIN_RAYS=262144 NNBVH_GRID_RAYS_PER_LANE=8" "NNBVH_WIDE_MINB=3" "NNBVH_WIDE_MINB=2"; do
  echo "== $v" >> $O/r02n_config3.txt
  env $v python tools/config3.py --cpu-rays 0 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l)
        print(d.get('bounce','all'), d['rays'], d['ms'], d['Mrays_s'])
" >> $O/r02n_config3.txt
done
cat $O/r02n_config3.txt
