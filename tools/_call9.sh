mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_t5_multigpu.py -m gpu -x -q > gpurun_out/c9_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c9_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29864 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c9_bench.json 2> gpurun_out/c9_bench.err; echo "rc=$?" >> gpurun_out/c9_bench.err
FTL_SPARSE_EMB_RS=0 timeout 900 $TR --master-port 29865 bench.py --gpus 2 --steps 10 --warmup 3 --no-ckpt --no-e2e --no-fsdp > gpurun_out/c9_bench_dense_emb.json 2> gpurun_out/c9_bench_dense.err
tail -6 gpurun_out/c9_pytest.log | cut -c1-300; cat gpurun_out/c9_bench.json gpurun_out/c9_bench_dense_emb.json
