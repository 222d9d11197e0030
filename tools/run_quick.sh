python -m pytest tests/test_gpu_parity.py -m gpu -q -k "coplanar or rough_bsdf_gradient_targets_parity" 2>&1 | grep -E "^E  |assert|Error|passed|failed" | head -40
