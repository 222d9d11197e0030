# Round-2 run C: GPU test suite on the rewritten wavefront (prefetched trace pipeline, no generate kernel, deferred NEE applied by shade)
set -x
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2c_pytest.log 2>&1; tail -40 gpurun_out/r2c_pytest.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-prb > gpurun_out/r2c_bench.log 2>&1; tail -1 gpurun_out/r2c_bench.log | cut -c1-3000
