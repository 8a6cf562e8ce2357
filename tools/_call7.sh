mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29841 tools/comm_roofline.py --mb 64 2> gpurun_out/c7_rf.err > gpurun_out/c7_roofline.jsonl
timeout 900 python -m pytest tests/test_t5_multigpu.py -m gpu -x -q -k "fsdp_tracks or stress" > gpurun_out/c7_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c7_pytest.log
timeout 900 $TR --master-port 29844 bench.py --gpus 2 --steps 10 --warmup 3 --no-ckpt --no-e2e > gpurun_out/c7_bench.json 2> gpurun_out/c7_bench.err; echo "rc=$?" >> gpurun_out/c7_bench.err
grep gather gpurun_out/c7_roofline.jsonl; tail -3 gpurun_out/c7_pytest.log; cat gpurun_out/c7_bench.json
