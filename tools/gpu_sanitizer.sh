#!/bin/bash
# compute-sanitizer passes over a slice of the parity tests: memcheck on the pair / team /
# full / chunked / resident paths, racecheck + synccheck on the kernels with mbarrier rings, named barriers and
# double-buffered shared memory.  Logs are kept under gpurun_out/ (summaries go to profiles/).
mkdir -p gpurun_out
SAN=/usr/local/cuda/bin/compute-sanitizer
SEL='dims_unrolled or component_counts or ragged_point or packed_noise_every or component_chunks or resident_fit or estep_full_ragged or staging'
timeout 2400 $SAN --tool memcheck --error-exitcode 86 --log-file gpurun_out/sanitizer_memcheck.log \
    python -m pytest tests/test_gpu_ops.py tests/test_layouts_and_chunks.py tests/test_team_kernels.py -m gpu -x -q -k "$SEL" > gpurun_out/sanitizer_memcheck_pytest.log 2>&1
echo "memcheck exit $?"; tail -3 gpurun_out/sanitizer_memcheck_pytest.log; tail -4 gpurun_out/sanitizer_memcheck.log
SEL2='(dims_unrolled and (10 or 16 or 9 or 12)) or (component_counts and (64 or 128 or 256 or 33)) or (packed_noise_every and (10 or 16 or 20 or 32))'
timeout 2400 $SAN --tool racecheck --racecheck-report all --error-exitcode 86 --log-file gpurun_out/sanitizer_racecheck.log \
    python -m pytest tests/test_gpu_ops.py tests/test_layouts_and_chunks.py -m gpu -x -q -k "$SEL2" > gpurun_out/sanitizer_racecheck_pytest.log 2>&1
echo "racecheck exit $?"; tail -3 gpurun_out/sanitizer_racecheck_pytest.log; tail -4 gpurun_out/sanitizer_racecheck.log
timeout 1200 $SAN --tool synccheck --error-exitcode 86 --log-file gpurun_out/sanitizer_synccheck.log \
    python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k "$SEL2" > gpurun_out/sanitizer_synccheck_pytest.log 2>&1
echo "synccheck exit $?"; tail -3 gpurun_out/sanitizer_synccheck_pytest.log; tail -4 gpurun_out/sanitizer_synccheck.log
