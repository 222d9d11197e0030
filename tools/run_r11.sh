python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu11.log; tail -2 gpurun_out/pytest_gpu11.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench11.log 2>&1; tail -1 gpurun_out/bench11.log | cut -c1-300
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench11_ref.log 2>&1; tail -1 gpurun_out/bench11_ref.log | cut -c1-600
