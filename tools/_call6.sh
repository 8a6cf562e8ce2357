mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for gw in 0 2 4 8; do
FTL_GATHER_WARPS=$gw timeout 300 $TR --master-port 2983$gw tools/comm_roofline.py --mb 64 2> gpurun_out/c6_rf_$gw.err | grep gemm_swiglu_gather >> gpurun_out/c6_gather_warps.jsonl
done
timeout 600 python -m pytest tests/test_native_engine.py tests/test_native_ops.py -m gpu -x -q > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c6_pytest.log
cat gpurun_out/c6_gather_warps.jsonl; tail -3 gpurun_out/c6_pytest.log
