mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c14_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c14_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c14_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/c14_smoke.log
timeout 700 python bench.py > gpurun_out/c14_bench.json 2> gpurun_out/c14_bench.err; echo "rc=$?" >> gpurun_out/c14_bench.err
tail -3 gpurun_out/c14_pytest.log | cut -c1-200; tail -2 gpurun_out/c14_smoke.log; cat gpurun_out/c14_bench.json
