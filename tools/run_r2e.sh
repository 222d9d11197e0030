# Round-2 run E: new tests (front door, optimisation loop, converged relMSE) + full bench line
set -x
timeout 1500 python -m pytest tests -m gpu -q --durations=6 > gpurun_out/r2e_pytest.log 2>&1; tail -25 gpurun_out/r2e_pytest.log
python examples/optimize_woodfloor.py --res 96 --iters 40 > gpurun_out/r2e_optimize.log 2>&1; tail -4 gpurun_out/r2e_optimize.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2e_bench.log 2>&1; tail -1 gpurun_out/r2e_bench.log | cut -c1-600
