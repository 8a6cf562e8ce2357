mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29921 bench.py --gpus 8 --steps 6 --warmup 3 --parallel fsdp --no-ckpt --no-e2e > gpurun_out/c19_bench_fsdp8_sg.json 2> gpurun_out/c19_bench_fsdp8.err; echo "rc=$?" >> gpurun_out/c19_bench_fsdp8.err
cat gpurun_out/c19_bench_fsdp8_sg.json | cut -c1-330; tail -3 gpurun_out/c19_bench_fsdp8.err | cut -c1-200
