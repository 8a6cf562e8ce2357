mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
FTL_FSDP_SHARD_GRADS=1 timeout 300 $TR --master-port 29911 tools/fsdp_check.py > gpurun_out/c17_fsdp_check4_sg.jsonl 2> gpurun_out/c17_fsdp_check4_sg.err; echo "rc=$?" >> gpurun_out/c17_fsdp_check4_sg.jsonl
FTL_FSDP_SHARD_GRADS=1 timeout 400 $TR --master-port 29912 bench.py --gpus 4 --steps 8 --warmup 3 --parallel fsdp --no-ckpt --no-e2e > gpurun_out/c17_bench_fsdp4_sg.json 2> gpurun_out/c17_bench_sg.err
timeout 400 $TR --master-port 29913 bench.py --gpus 4 --steps 8 --warmup 3 --parallel fsdp --no-ckpt --no-e2e > gpurun_out/c17_bench_fsdp4_def.json 2> gpurun_out/c17_bench_def.err
cat gpurun_out/c17_fsdp_check4_sg.jsonl | cut -c1-300; cat gpurun_out/c17_bench_fsdp4_sg.json gpurun_out/c17_bench_fsdp4_def.json | cut -c1-330
