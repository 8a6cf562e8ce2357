mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29871 tools/comm_stress.py --epochs 200 --gemms 10 > gpurun_out/c9b_stress.out 2> gpurun_out/c9b_stress.err; echo "rc=$?" >> gpurun_out/c9b_stress.out
grep -v "NCCL\|^$" gpurun_out/c9b_stress.err | grep -B2 -A25 "Traceback" | head -60; cat gpurun_out/c9b_stress.out
