cd $GRAFT_REPO_ROOT
O=gpurun_out
: > $O/r02zj_sweep.txt
CF=""
for cd in 4,1 4,2 3,2 5,2 5,1 4,3 6,1; do c=${cd%,*}; d=${cd#*,}; CF="$CF NNBVH_KERNEL=3,NNBVH_BIN_CELL_BITS=$c,NNBVH_BIN_DIR_BITS=$d"; done
for sc in bathroom crown; do
echo "$sc" >> $O/r02zj_sweep.txt; python tools/sweep.py $sc 8388608 NNBVH_KERNEL=3 $CF 2>&1 | tail -8 >> $O/r02zj_sweep.txt
done
cat $O/r02zj_sweep.txt
