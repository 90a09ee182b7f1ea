# one GPU round: tests, smoke, bench (both arms), launch list, ncu --set full of the named kernels.  TAG=r1c bash scripts/gpu_round.sh
TAG=${TAG:-r1x}; KERNELS=${KERNELS:-"k_cigar k_extend2 k_seed2"}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${TAG}_smoke.log
python bench.py --steps 6 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 --pairs 200000 --no-cpu-baseline > gpurun_out/${TAG}_ncu1.log 2>&1
for k in $KERNELS; do
ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o gpurun_out/${TAG}_$k -f python bench.py --steps 1 --warmup 1 --pairs ${NCU_PAIRS:-200000} --no-cpu-baseline > gpurun_out/${TAG}_ncu_$k.log 2>&1
done
tail -3 gpurun_out/${TAG}_pytest.log; tail -2 gpurun_out/${TAG}_smoke.log; cat gpurun_out/${TAG}_bench.json; cat gpurun_out/${TAG}_bench_ref.json; tail -5 gpurun_out/${TAG}_bench.err
