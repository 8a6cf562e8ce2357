mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/c1_smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
timeout 120 python tools/norm_time.py > gpurun_out/c1_norm.json 2> gpurun_out/c1_norm.err
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/c1_bench_ours.json 2> gpurun_out/c1_bench_ours.err; echo "rc=$?" >> gpurun_out/c1_bench_ours.err
FTL_FUSED_NORM=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-ckpt --no-e2e > gpurun_out/c1_bench_ours_nofusednorm.json 2> gpurun_out/c1_bench_ours_nfn.err
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/c1_bench_ref.json 2> gpurun_out/c1_bench_ref.err; echo "rc=$?" >> gpurun_out/c1_bench_ref.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:rmsnorm_fwd -c 4 -o gpurun_out/c1_norm_ncu python tools/norm_time.py > gpurun_out/c1_ncu.log 2>&1
tail -5 gpurun_out/c1_pytest.log; cat gpurun_out/c1_norm.json gpurun_out/c1_bench_ours.json gpurun_out/c1_bench_ours_nofusednorm.json gpurun_out/c1_bench_ref.json
