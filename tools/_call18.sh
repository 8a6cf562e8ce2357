mkdir -p gpurun_out
timeout 1700 python -m pytest tests/test_t5_multigpu.py -m gpu -x -q > gpurun_out/c18_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c18_pytest.log
tail -6 gpurun_out/c18_pytest.log | cut -c1-300
