#!/bin/bash
# tools/sanitize.sh — SURVEY §5 row 2 / VERDICT r1 item 6.  Run on the GPU box (gpurun):
#   compute-sanitizer --tool racecheck / memcheck over tools/sanitize_target.py (the warp-shared triangle queue of
#   trace_wq folds hits with shared-memory atomics between __syncwarp()s: the code racecheck exists for);
#   then (CPU) an AddressSanitizer + UBSan build of the host BVH builder run through tests/test_host_builder.py.
# Logs land in gpurun_out/sanitize_*.log; the summaries kept in the repo are profiles/r2_sanitize_*.txt.
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
for w in render deep masked; do
  timeout 900 $CS --tool racecheck --racecheck-report all --print-limit 20 python tools/sanitize_target.py $w \
      > gpurun_out/sanitize_racecheck_$w.log 2>&1; echo "exit $?" >> gpurun_out/sanitize_racecheck_$w.log
  timeout 900 $CS --tool memcheck --print-limit 20 python tools/sanitize_target.py $w \
      > gpurun_out/sanitize_memcheck_$w.log 2>&1; echo "exit $?" >> gpurun_out/sanitize_memcheck_$w.log
done
grep -H -E "RACECHECK SUMMARY|ERROR SUMMARY|^exit" gpurun_out/sanitize_*check_*.log
# host side: ASan + UBSan build of the builder + the __host__ traversal core
( cd smallpt-enoki-optix_b200 && /usr/bin/g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fno-omit-frame-pointer \
    -ffp-contract=off -fopenmp -fPIC -shared -Wall -Wno-unknown-pragmas -I/usr/local/cuda/include \
    -o build/libwpt_hosttest_asan.so csrc/hosttest.cpp csrc/bvh_build.cpp ) && \
WPT_HOSTTEST_LIB=$PWD/smallpt-enoki-optix_b200/build/libwpt_hosttest_asan.so \
LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libubsan.so)" \
ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1 \
  timeout 900 python -m pytest tests/test_host_builder.py -q -x -p no:cacheprovider > gpurun_out/sanitize_host_asan.log 2>&1
echo "host asan exit $?" >> gpurun_out/sanitize_host_asan.log; tail -3 gpurun_out/sanitize_host_asan.log
