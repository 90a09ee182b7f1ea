# ncu passes: launch list restricted by regex + one full capture per kernel.  TAG=r2d KREGEX=k_ext KERNELS="k_ext_dp" bash scripts/gpu_ncu.sh
TAG=${TAG:-r2x}; KREGEX=${KREGEX:-.}; PAIRS=${PAIRS:-1000000}; CFG=${CFG:-haplotag}
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:$KREGEX -c ${NLAUNCH:-400} --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --config $CFG --steps 1 --warmup 1 --pairs $PAIRS --no-cpu-baseline > gpurun_out/${TAG}_ncu1.log 2>&1
python - <<PY
import csv,collections
rows=[r for r in csv.reader(open("gpurun_out/${TAG}_launches.csv")) if len(r)>10 and r[0].isdigit()]
agg=collections.OrderedDict()
for r in rows:
    name=r[4].split('(')[0]; v=float(r[-1].replace(',','')); u=r[-2]
    v = v/1e3 if u in('nsecond','ns') else v*1e3 if u in ('msecond','ms') else v
    a=agg.setdefault(name,[0,0.0]); a[0]+=1; a[1]+=v
for k,(n,t) in agg.items(): print("%-60s n=%4d total=%10.1f us"%(k[:60],n,t))
PY
for k in $KERNELS; do
ncu --set full --clock-control none --import-source on -k regex:$k -s ${SKIP:-0} -c 1 -o gpurun_out/${TAG}_$k -f python bench.py --config $CFG --steps 1 --warmup 1 --pairs $PAIRS --no-cpu-baseline > gpurun_out/${TAG}_ncu_$k.log 2>&1
done
