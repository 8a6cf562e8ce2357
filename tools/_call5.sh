mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_t5_multigpu.py -m gpu -x -q -k "fsdp_tracks or stress" > gpurun_out/c5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c5_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29822 tools/comm_roofline.py --mb 256 > gpurun_out/c5_roofline.jsonl 2> gpurun_out/c5_roofline.err; echo "rc=$?" >> gpurun_out/c5_roofline.err
timeout 600 $TR --master-port 29825 tools/step_kernels_time.py --parallel fsdp > gpurun_out/c5_step_kernels_fsdp2.md 2> gpurun_out/c5_sk_fsdp.err
timeout 600 $TR --master-port 29826 tools/step_kernels_time.py --parallel dp > gpurun_out/c5_step_kernels_dp2.md 2> gpurun_out/c5_sk_dp.err
timeout 900 $TR --master-port 29824 bench.py --gpus 2 --steps 10 --warmup 3 --no-ckpt --no-e2e > gpurun_out/c5_bench.json 2> gpurun_out/c5_bench.err; echo "rc=$?" >> gpurun_out/c5_bench.err
tail -4 gpurun_out/c5_pytest.log; cat gpurun_out/c5_roofline.jsonl gpurun_out/c5_bench.json; head -30 gpurun_out/c5_step_kernels_fsdp2.md
