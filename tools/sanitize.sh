#!/bin/bash
# compute-sanitizer over the small-template GPU tests (ONE tool per gpurun call, B200_PROFILING.md).
#   bash tools/sanitize.sh memcheck|racecheck|synccheck|initcheck  -> gpurun_out/r2_sanitizer_<tool>.log
tool=${1:-memcheck}
TESTS="tests/test_gpu_native.py tests/test_reference_golden.py tests/test_reference_dimers.py tests/test_mixture_join.py tests/test_zz_no_alloc.py"
# the plain run first: a program that faults must not go under the tool
python -m pytest $TESTS -m gpu -x -q > gpurun_out/r2_sanitizer_plain.log 2>&1 || { tail -5 gpurun_out/r2_sanitizer_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $tool --error-exitcode 99 --log-file gpurun_out/r2_sanitizer_$tool.log \
  python -m pytest $TESTS -m gpu -x -q > gpurun_out/r2_sanitizer_${tool}_pytest.log 2>&1
echo "rc=$?"
tail -3 gpurun_out/r2_sanitizer_${tool}_pytest.log
grep -c "Invalid\|Race\|hazard\|Uninitialized" gpurun_out/r2_sanitizer_$tool.log
tail -4 gpurun_out/r2_sanitizer_$tool.log
