#!/bin/bash
# usage: gpu_sanitize.sh memcheck|racecheck : compute-sanitizer over smoke(), the fuzzed decoder and one small-batch case
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool $TOOL --log-file gpurun_out/sanitizer_$TOOL.log --print-limit 50 \
  python -m pytest "tests/test_fuzz_gpu.py" "tests/test_batch_gpu.py::test_batch_size_invariance[40-128]" "tests/test_batch_gpu.py::test_batch_size_invariance[97-96]" tests/test_zz_no_alloc.py tests/test_encode_gpu.py -m gpu -q -x > gpurun_out/sanitizer_${TOOL}_pytest.log 2>&1
echo "pytest under $TOOL rc $?"; tail -n 3 gpurun_out/sanitizer_${TOOL}_pytest.log
timeout 600 compute-sanitizer --tool $TOOL --log-file gpurun_out/sanitizer_${TOOL}_smoke.log --print-limit 50 python __graft_entry__.py --smoke > gpurun_out/sanitizer_${TOOL}_smoke.out 2>&1
echo "smoke under $TOOL rc $?"; tail -n 2 gpurun_out/sanitizer_${TOOL}_smoke.out
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Invalid|hazard" gpurun_out/sanitizer_$TOOL.log gpurun_out/sanitizer_${TOOL}_smoke.log | sort | uniq -c | head -20
