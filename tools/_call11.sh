mkdir -p gpurun_out
which compute-sanitizer > gpurun_out/c11_sanitize.log 2>&1
timeout 120 compute-sanitizer --tool memcheck --print-limit 5 python tools/gemm_check.py nt_small > gpurun_out/c11_first.log 2>&1; echo "first rc=$?" >> gpurun_out/c11_sanitize.log
head -c 3000 gpurun_out/c11_first.log >> gpurun_out/c11_sanitize.log
timeout 900 bash tools/sanitize.sh memcheck >> gpurun_out/c11_sanitize.log 2>&1
tail -40 gpurun_out/c11_sanitize.log | cut -c1-220
