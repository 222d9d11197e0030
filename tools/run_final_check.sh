python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py > gpurun_out/final2_bench.log 2>&1; tail -1 gpurun_out/final2_bench.log | cut -c1-250
