"""Generates tests/golden/c1_oracle_16384spp.npz: the CPU oracle's converged image of BASELINE config C1 (mi.cornell_box()
256 x 256, path.py loop, max_depth 8, rr_depth 4) at 16 384 spp, and the relMSE of the oracle's own 4 096-spp render (another
seed) against it -- the Monte-Carlo noise floor the CUDA path's 4 096-spp relMSE is compared with
(tests/test_gpu_configs.py, bench.py `parity`).  About 20 minutes on 8 cores:  python tests/golden/make_c1_converged.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from mitsuba3_experiments_b200.scene import SceneDesc  # noqa: E402
from mitsuba3_experiments_b200.scenes import cornell_box  # noqa: E402
from oracle.oracle_py import OracleScene  # noqa: E402

orc = OracleScene(SceneDesc(cornell_box(256, 256, 16)))
kw = dict(max_depth=8, rr_depth=4, variant=0)
ref, _ = orc.render(seed=12345, spp=16384, **kw)
low, _ = orc.render(seed=777, spp=4096, **kw)
relmse = float(np.mean((low - ref) ** 2 / (ref ** 2 + 1e-2)))
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "c1_oracle_16384spp.npz"), image=ref.astype(np.float32),
                    relmse_oracle_4096=np.float64(relmse), seed=12345, spp=16384)
print("relMSE(oracle 4096 spp vs oracle 16384 spp) =", relmse)
