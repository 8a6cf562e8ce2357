mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_t5_multigpu.py tests/test_native_engine.py -m gpu -x -q -k "nvls or fsdp or nonfinite or stress" > gpurun_out/c4_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c4_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29822 tools/comm_roofline.py --mb 256 --nvls > gpurun_out/c4_roofline_nvls.jsonl 2> gpurun_out/c4_roofline_nvls.err; echo "rc=$?" >> gpurun_out/c4_roofline_nvls.err
timeout 900 $TR --master-port 29824 bench.py --gpus 2 --steps 10 --warmup 3 --no-ckpt --no-e2e > gpurun_out/c4_bench.json 2> gpurun_out/c4_bench.err; echo "rc=$?" >> gpurun_out/c4_bench.err
FTL_COLLECTIVE=nvls timeout 900 $TR --master-port 29823 bench.py --gpus 2 --steps 10 --warmup 3 --no-ckpt --no-e2e > gpurun_out/c4_bench_nvls.json 2> gpurun_out/c4_bench_nvls.err; echo "rc=$?" >> gpurun_out/c4_bench_nvls.err
tail -8 gpurun_out/c4_pytest.log; cat gpurun_out/c4_roofline_nvls.jsonl gpurun_out/c4_bench.json gpurun_out/c4_bench_nvls.json; tail -n 4 gpurun_out/c4_bench.err gpurun_out/c4_bench_nvls.err
