mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29891 tools/dp_check.py > gpurun_out/c13_dp_check4.jsonl 2> gpurun_out/c13_dp_check4.err; echo "rc=$?" >> gpurun_out/c13_dp_check4.jsonl
timeout 300 $TR --master-port 29892 tools/fsdp_check.py > gpurun_out/c13_fsdp_check4.jsonl 2> gpurun_out/c13_fsdp_check4.err; echo "rc=$?" >> gpurun_out/c13_fsdp_check4.jsonl
timeout 500 $TR --master-port 29893 bench.py --gpus 4 --steps 8 --warmup 3 --no-ckpt > gpurun_out/c13_bench4.json 2> gpurun_out/c13_bench4.err; echo "rc=$?" >> gpurun_out/c13_bench4.err
cat gpurun_out/c13_dp_check4.jsonl gpurun_out/c13_fsdp_check4.jsonl | cut -c1-500; cat gpurun_out/c13_bench4.json
