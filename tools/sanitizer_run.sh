cd /root/repo
export DFQ_WATCHDOG_S=900
echo "== memcheck: smoke()"; timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -6
echo "== memcheck: q8 pointwise (tensor-map TMA + tcgen05) + depthwise + stem/add/mean, cle pairs (resident)"; timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_q8.py -q -m gpu -k "pointwise or depthwise or stem" -x 2>&1 | tail -6
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_pipeline.py -q -m gpu -k "pairs_against_oracle or long_chain" -x 2>&1 | tail -6
echo "== racecheck: smoke()"; timeout 900 compute-sanitizer --tool racecheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -8
echo "== racecheck: cle3 + q8 pointwise small"; timeout 900 compute-sanitizer --tool racecheck --error-exitcode 7 python -m pytest tests/test_gpu_pipeline.py -q -m gpu -k "pairs_against_oracle and resident" -x 2>&1 | tail -6
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 7 python -m pytest "tests/test_gpu_q8.py::test_pointwise_q8_against_integer_restatement[1000-16-96-0-1-False]" -q -m gpu -x 2>&1 | tail -6
