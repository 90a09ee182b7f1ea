export ARACHNE_SEED_IMPL=3
for t in ${THREADS:-512 768 1024}; do
  export ARACHNE_SEED3_THREADS=$t
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('threads $t', {k: round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if k in ('seed','total')})"
done
