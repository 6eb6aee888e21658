#!/usr/bin/env python
"""Times the tensor-pipe DFT-pass experiment (csrc/tc_dft.cu) against the scalar butterflies on the same data:
  python scripts/tc_experiment.py [rows=1<<20] [reps=3]  ->  one JSON line (also the numbers' error vs float64 at reps=1)
rows x 32 complex points per launch; reps = passes run back to back on every tile in shared memory / TMEM (the
production transform has three).  Run on the GPU box; profiles/r2_tc_experiment.json keeps the result."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from narrow_band_least_squares_b200.engine import Engine  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
rng = np.random.default_rng(1)
x = (rng.standard_normal((rows, 32)) + 1j * rng.standard_normal((rows, 32))).astype(np.complex64)
eng = Engine(0)
n = np.arange(32)
dft = np.exp(2j * np.pi * np.outer(n, n) / 32)
w1 = np.exp(2j * np.pi * (np.arange(4096) % 256) / 8192)
ref = (x[:4096].astype(np.complex128) @ dft.T) * (w1[:, None] ** n[None, :])
out = {"rows": rows, "points_per_launch": rows * 32, "passes_per_launch": reps}
for name, mode in (("scalar_butterflies", 0), ("tcgen05_fp16_single_product", 1),
                   ("tcgen05_fp16_split_three_products", 2)):
    y1, _ = eng.experiment_dft_pass(x[:4096], mode, reps=1, iters=1)
    err = float(np.max(np.abs(y1 - ref)) / np.max(np.abs(ref)))
    _, ms = eng.experiment_dft_pass(x, mode, reps=reps, iters=5)
    _, ms1 = eng.experiment_dft_pass(x, mode, reps=1, iters=5)
    out[name] = {"ms": ms, "ms_one_pass_launch": ms1,
                 "ns_per_point_pass": (ms - ms1) * 1e6 / (rows * 32 * max(reps - 1, 1)), "rel_err_one_pass": err}
print(json.dumps(out))
