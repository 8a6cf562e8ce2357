mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_t5_multigpu.py -m gpu -x -q -k "stress or vmm or nvls or chain" > gpurun_out/c9c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c9c_pytest.log
tail -6 gpurun_out/c9c_pytest.log | cut -c1-300
