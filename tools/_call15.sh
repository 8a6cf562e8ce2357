mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_native_ops.py tests/test_native_engine.py -m gpu -x -q -k "chunked_head or autograd or gemm_tc" > gpurun_out/c15_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c15_pytest.log
FTL_FUSED_HEAD=1 timeout 500 python bench.py --steps 8 --warmup 3 --no-ckpt --no-e2e > gpurun_out/c15_bench_fused_head.json 2> gpurun_out/c15_bench_fused_head.err; echo "rc=$?" >> gpurun_out/c15_bench_fused_head.err
timeout 500 python bench.py --steps 8 --warmup 3 --no-ckpt --no-e2e > gpurun_out/c15_bench_default.json 2> gpurun_out/c15_bench_default.err
FTL_FUSED_HEAD=1 timeout 500 python bench.py --steps 4 --warmup 3 --seq 8192 --no-ckpt --no-e2e > gpurun_out/c15_bench_fused_head_8192.json 2> gpurun_out/c15_bench_fused_head_8192.err
tail -12 gpurun_out/c15_pytest.log | cut -c1-300; cat gpurun_out/c15_bench_fused_head.json gpurun_out/c15_bench_default.json gpurun_out/c15_bench_fused_head_8192.json | cut -c1-400; tail -3 gpurun_out/c15_bench_fused_head.err | cut -c1-300
