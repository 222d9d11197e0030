# Round-2 run B: full GPU test suite on the round-1 kernels with the new config-scale tests (baseline for the kernel rewrite)
set -x
timeout 1500 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/r2b_pytest.log 2>&1; tail -30 gpurun_out/r2b_pytest.log
