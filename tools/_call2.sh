mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/c2_topo.txt 2>&1
timeout 1500 python -m pytest tests/test_t5_multigpu.py -m gpu -x -q > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c2_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29811 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c2_bench_ours.json 2> gpurun_out/c2_bench_ours.err; echo "rc=$?" >> gpurun_out/c2_bench_ours.err
NCCL_DEBUG=INFO timeout 1200 $TR --master-port 29812 bench.py --impl reference --gpus 2 --steps 10 --warmup 3 > gpurun_out/c2_bench_ref.json 2> gpurun_out/c2_bench_ref.err; echo "rc=$?" >> gpurun_out/c2_bench_ref.err
grep -m3 -i "nvls" gpurun_out/c2_bench_ref.err > gpurun_out/c2_nccl_nvls.txt
tail -15 gpurun_out/c2_pytest.log; cat gpurun_out/c2_bench_ours.json gpurun_out/c2_bench_ref.json; tail -5 gpurun_out/c2_bench_ours.err; tail -5 gpurun_out/c2_bench_ref.err
