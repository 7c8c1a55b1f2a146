timeout 120 python tools/sanitize_small.py > gpurun_out/r1p_sanitize_plain.log 2>&1 && timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_small.py > gpurun_out/r1p_memcheck.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/r1p_sanitize_plain.log; tail -8 gpurun_out/r1p_memcheck.log
