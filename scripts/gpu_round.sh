set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1b_pytest.log
python bench.py --steps 6 --warmup 3 > gpurun_out/r1b_bench.json 2> gpurun_out/r1b_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1b_bench_ref.json 2>> gpurun_out/r1b_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1b_launches.csv python bench.py --steps 2 --warmup 1 --pairs 200000 --no-cpu-baseline > gpurun_out/r1b_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_seed2 -s 1 -c 1 -o gpurun_out/r1b_seed -f python bench.py --steps 1 --warmup 1 --pairs 200000 --no-cpu-baseline > gpurun_out/r1b_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_extend2 -s 1 -c 1 -o gpurun_out/r1b_extend -f python bench.py --steps 1 --warmup 1 --pairs 200000 --no-cpu-baseline > gpurun_out/r1b_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_cigar -s 1 -c 1 -o gpurun_out/r1b_cigar -f python bench.py --steps 1 --warmup 1 --pairs 200000 --no-cpu-baseline > gpurun_out/r1b_ncu4.log 2>&1
nproc; free -g | head -2
tail -3 gpurun_out/r1b_pytest.log; cat gpurun_out/r1b_bench.json; cat gpurun_out/r1b_bench_ref.json
