#!/bin/bash
# compute-sanitizer over the GPU parity tests (SURVEY §5 "race detection"): memcheck on factor / map / ba, racecheck on the kernels with
# hand-placed barriers and shared-memory hand-offs (fused Hessian, batched LM solve, blocked LDL^T).  Summaries -> gpurun_out/.
export PYTHONDONTWRITEBYTECODE=1
CS=/usr/local/cuda/bin/compute-sanitizer
timeout 700 $CS --tool memcheck --target-processes all --print-limit 20 python -m pytest tests/test_factor_gpu.py tests/test_ba_gpu.py tests/test_map_gpu.py tests/test_hba_gpu.py -x -q -m gpu -p no:cacheprovider > gpurun_out/sanitizer_memcheck.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/sanitizer_memcheck.log
timeout 700 $CS --tool racecheck --target-processes all --print-limit 20 python -m pytest tests/test_factor_gpu.py tests/test_ba_gpu.py -x -q -m gpu -p no:cacheprovider > gpurun_out/sanitizer_racecheck.log 2>&1
echo "racecheck rc=$?" >> gpurun_out/sanitizer_racecheck.log
tail -5 gpurun_out/sanitizer_memcheck.log; tail -5 gpurun_out/sanitizer_racecheck.log
