"""The job scripts/sanitize.sh runs under compute-sanitizer: config 1 (OLS) and a short LTS array, every stage."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from narrow_band_least_squares_b200.configs import config_plan  # noqa: E402
from narrow_band_least_squares_b200.engine import Engine, Results  # noqa: E402
from narrow_band_least_squares_b200.synth import config_record, plane_wave_record  # noqa: E402
from narrow_band_least_squares_b200.narrow_band_least_squares import build_plan  # noqa: E402

hours = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
eng = Engine(0)
rec = config_record("c1", hours=hours)
plan = config_plan("c1", rec.npts, rec.rij)
res = eng.run(plan, rec.data, Results(plan, pinned=False))
print("c1:", plan.nwin, "estimates, median vel", float(np.nanmedian(res.vel[:plan.nwin])))
rec = plane_wave_record(seed=7, nchans=6, npts=int(6000 * max(hours, 0.2)), fs=50.0, aperture_km=0.4, flip_channel=3)
fl = np.logspace(np.log10(0.5), np.log10(5.0), 4)
plan = build_plan(6, rec.npts, rec.fs, rec.rij, [20.0] * 3, 0.5, 0.75, 3, fl, "log", "butter", 2, 0.01, quiet=True)
res = eng.run(plan, rec.data, Results(plan, pinned=False))
print("lts:", plan.nwin, "estimates, flagged windows", int((res.weights[:plan.nwin] == 0).any(axis=1).sum()))
eng.close()
