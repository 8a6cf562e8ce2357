mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/c8_topo.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_t5_multigpu.py -m gpu -x -q -k "tracks and 8" > gpurun_out/c8_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c8_pytest.log
timeout 700 $TR --master-port 29851 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/c8_bench_ours.json 2> gpurun_out/c8_bench_ours.err; echo "rc=$?" >> gpurun_out/c8_bench_ours.err
timeout 900 $TR --master-port 29852 bench.py --impl reference --gpus 8 --steps 10 --warmup 3 > gpurun_out/c8_bench_ref.json 2> gpurun_out/c8_bench_ref.err; echo "rc=$?" >> gpurun_out/c8_bench_ref.err
timeout 300 $TR --master-port 29853 tools/comm_roofline.py --mb 256 --nvls > gpurun_out/c8_roofline.jsonl 2> gpurun_out/c8_roofline.err; echo "rc=$?" >> gpurun_out/c8_roofline.err
timeout 300 $TR --master-port 29854 tools/step_kernels_time.py --parallel dp > gpurun_out/c8_step_kernels_dp8.md 2> gpurun_out/c8_sk_dp.err
timeout 300 $TR --master-port 29855 tools/step_kernels_time.py --parallel fsdp > gpurun_out/c8_step_kernels_fsdp8.md 2> gpurun_out/c8_sk_fsdp.err
FTL_COLLECTIVE=nvls timeout 400 $TR --master-port 29856 bench.py --gpus 8 --steps 10 --warmup 3 --no-ckpt --no-e2e > gpurun_out/c8_bench_nvls.json 2> gpurun_out/c8_bench_nvls.err; echo "rc=$?" >> gpurun_out/c8_bench_nvls.err
tail -4 gpurun_out/c8_pytest.log; cat gpurun_out/c8_bench_ours.json gpurun_out/c8_bench_ref.json gpurun_out/c8_bench_nvls.json; cat gpurun_out/c8_roofline.jsonl | cut -c1-330
