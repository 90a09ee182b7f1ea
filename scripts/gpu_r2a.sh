# round-2 first GPU pass: tests, the three workload configs, sanitizer
TAG=${TAG:-r2a}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
for cfg in haplotag stlfr tellseq; do
  timeout 900 python bench.py --config $cfg --steps 4 --warmup 3 > gpurun_out/${TAG}_bench_$cfg.json 2> gpurun_out/${TAG}_bench_$cfg.err; echo "bench $cfg rc=$?"
  cat gpurun_out/${TAG}_bench_$cfg.json; tail -4 gpurun_out/${TAG}_bench_$cfg.err
done
TOOLS="memcheck racecheck" SAN_TIMEOUT=420 TAG=$TAG bash scripts/sanitize.sh
