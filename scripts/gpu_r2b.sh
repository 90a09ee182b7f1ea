TAG=${TAG:-r2b}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
for cfg in ${CFGS:-stlfr}; do
  timeout 900 python bench.py --config $cfg --steps 4 --warmup 3 > gpurun_out/${TAG}_bench_$cfg.json 2> gpurun_out/${TAG}_bench_$cfg.err; echo "bench $cfg rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_$cfg.json"))
print("$cfg", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "stages", {k: round(v,2) for k,v in d["roofline"]["stage_ms_per_step"].items()}, "parity", d.get("parity_sample",{}).get("ok"), d.get("parity_sample",{}).get("placement_mismatches"))
PY
  tail -2 gpurun_out/${TAG}_bench_$cfg.err
done
