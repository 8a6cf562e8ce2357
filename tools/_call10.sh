mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c10_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c10_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c10_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/c10_smoke.log
timeout 700 python bench.py --steps 10 --warmup 3 > gpurun_out/c10_bench.json 2> gpurun_out/c10_bench.err; echo "rc=$?" >> gpurun_out/c10_bench.err
timeout 700 python bench.py --steps 5 --warmup 3 --seq 8192 --no-ckpt > gpurun_out/c10_bench_seq8192.json 2> gpurun_out/c10_bench_seq8192.err; echo "rc=$?" >> gpurun_out/c10_bench_seq8192.err
timeout 900 python bench.py --impl reference --steps 5 --warmup 3 --seq 8192 --no-ckpt > gpurun_out/c10_bench_ref_seq8192.json 2> gpurun_out/c10_bench_ref_seq8192.err; echo "rc=$?" >> gpurun_out/c10_bench_ref_seq8192.err
timeout 120 python tools/comm_single_gpu_case.py > gpurun_out/c10_comm_case.json 2> gpurun_out/c10_comm_case.err && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"dp_reduce_scatter|dp_adamw_allgather|fsdp_gather|gemm_tc2_kernel" -o gpurun_out/c10_comm_ncu python tools/comm_single_gpu_case.py --once > gpurun_out/c10_ncu.log 2>&1
tail -4 gpurun_out/c10_pytest.log | cut -c1-300; cat gpurun_out/c10_smoke.log | tail -2; cat gpurun_out/c10_bench.json gpurun_out/c10_bench_seq8192.json gpurun_out/c10_bench_ref_seq8192.json gpurun_out/c10_comm_case.json; tail -3 gpurun_out/c10_ncu.log
