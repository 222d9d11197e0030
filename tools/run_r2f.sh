python examples/optimize_woodfloor.py --res 96 2>/dev/null | tail -1
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "optimisation or front_door or autograd" 2>&1 | tail -3
