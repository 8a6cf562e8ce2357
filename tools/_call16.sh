mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_t5_multigpu.py -m gpu -x -q -k "sharded_gradients or fsdp_tracks" > gpurun_out/c16_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c16_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
FTL_FSDP_SHARD_GRADS=1 timeout 600 $TR --master-port 29901 bench.py --gpus 2 --steps 8 --warmup 3 --parallel fsdp --no-ckpt --no-e2e > gpurun_out/c16_bench_fsdp2_shardgrads.json 2> gpurun_out/c16_bench_sg.err; echo "rc=$?" >> gpurun_out/c16_bench_sg.err
timeout 600 $TR --master-port 29902 bench.py --gpus 2 --steps 8 --warmup 3 --parallel fsdp --no-ckpt --no-e2e > gpurun_out/c16_bench_fsdp2_default.json 2> gpurun_out/c16_bench_def.err
tail -8 gpurun_out/c16_pytest.log | cut -c1-300; cat gpurun_out/c16_bench_fsdp2_shardgrads.json gpurun_out/c16_bench_fsdp2_default.json | cut -c1-330; tail -5 gpurun_out/c16_bench_sg.err | cut -c1-300
