# multi-GPU bench line: N=$1 CONFIG=$2 [env...]   (run under gpurun --gpus N)
N=$1; CFG=$2; TAG=${TAG:-r2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --config $CFG --steps ${STEPS:-4} --warmup 3 > gpurun_out/${TAG}_bench_${CFG}_${N}gpu${SUFFIX}.json 2> gpurun_out/${TAG}_bench_${CFG}_${N}gpu${SUFFIX}.err
echo "rc=$?"; python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_${CFG}_${N}gpu${SUFFIX}.json"))
print("$CFG N=$N", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"],1), {k: round(v,1) for k,v in d["roofline"]["stage_ms_per_step"].items()})
PY
tail -2 gpurun_out/${TAG}_bench_${CFG}_${N}gpu${SUFFIX}.err
