# compute-sanitizer over a subset of the -m gpu suite on the small test references (SURVEY section 5 plan).
# TAG=r2a bash scripts/sanitize.sh   -> gpurun_out/${TAG}_sanitize_<tool>.log (+ a one-line summary per tool)
TAG=${TAG:-r2x}
TESTS=${TESTS:-"tests/test_gpu_parity.py::test_ksw_extend_parity tests/test_gpu_parity.py::test_rescue_and_cigar_parity tests/test_gpu_parity.py::test_place_parity tests/test_gpu_parity.py::test_config_flavours_all_stages tests/test_gpu_parity.py::test_seed_intervals_every_shortcut tests/test_gpu_parity.py::test_edge_inputs_all_stages"}
for tool in ${TOOLS:-memcheck racecheck initcheck}; do
  log=gpurun_out/${TAG}_sanitize_${tool}.log
  timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool $tool --target-processes all --print-limit 20 --log-file ${log}.raw \
      python -m pytest $TESTS -x -q -m gpu > ${log}.pytest 2>&1
  echo "rc=$?" >> ${log}.pytest
  { echo "== $tool =="; tail -3 ${log}.pytest; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|LEAK SUMMARY" ${log}.raw | sort | uniq -c; grep -E "Invalid|Race reported|Uninitialized|hazard" ${log}.raw | sed 's/0x[0-9a-f]*/0x../g' | sort | uniq -c | sort -rn | head -30; } > $log
  head -c 3000000 ${log}.raw > ${log}.raw.head && rm -f ${log}.raw
  cat $log
done
