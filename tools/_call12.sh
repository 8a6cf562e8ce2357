mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29881 tools/fsdp_check.py > gpurun_out/c12_fsdp_check8.jsonl 2> gpurun_out/c12_fsdp_check8.err; echo "rc=$?" >> gpurun_out/c12_fsdp_check8.jsonl
timeout 200 $TR --master-port 29883 tools/comm_stress.py --epochs 1000 --gemms 30 > gpurun_out/c12_stress8.jsonl 2> gpurun_out/c12_stress8.err; echo "rc=$?" >> gpurun_out/c12_stress8.jsonl
timeout 600 bash tools/ft_gpu_scenarios.sh fsdp8 > gpurun_out/c12_ft_fsdp8.txt 2>&1
timeout 500 $TR --master-port 29884 bench.py --gpus 8 --seq 8192 --steps 5 --warmup 3 --no-ckpt > gpurun_out/c12_bench_seq8192_8gpu.json 2> gpurun_out/c12_bench_seq8192.err; echo "rc=$?" >> gpurun_out/c12_bench_seq8192.err
timeout 300 $TR --master-port 29885 bench.py --gpus 8 --steps 10 --warmup 3 --no-ckpt --no-fsdp --no-e2e > gpurun_out/c12_bench_dp8_sparse.json 2> gpurun_out/c12_bench_dp8.err
cat gpurun_out/c12_fsdp_check8.jsonl gpurun_out/c12_stress8.jsonl | cut -c1-400; cat gpurun_out/c12_ft_fsdp8.txt | cut -c1-200; cat gpurun_out/c12_bench_seq8192_8gpu.json gpurun_out/c12_bench_dp8_sparse.json
