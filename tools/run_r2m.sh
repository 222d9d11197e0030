# Round-2 run M: full GPU suite + smoke + bench + reference arm on the final state
set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2m_smoke.log 2>&1; tail -1 gpurun_out/r2m_smoke.log
timeout 1500 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/r2m_pytest.log 2>&1; tail -12 gpurun_out/r2m_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2m_bench.log 2>&1; tail -1 gpurun_out/r2m_bench.log | cut -c1-300
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2m_bench_ref.log 2>&1; tail -1 gpurun_out/r2m_bench_ref.log | cut -c1-300
