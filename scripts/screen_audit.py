#!/usr/bin/env python
"""FP32-screen error measured at EVERY searched lag (NBLS_OPT_SCREEN_AUDIT) on the plane-wave c3 bands and on the
adversarial records of tests/test_gpu_wide.py; prints one JSON object (kept as profiles/r2_screen_audit.json).
The candidate band of the screen is XF_TAU32 = 2e-5: the true arg-max is among the candidates while this error stays
below 1e-5.  Run on the GPU box."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from narrow_band_least_squares_b200 import _capi  # noqa: E402
from narrow_band_least_squares_b200.configs import config_job  # noqa: E402
from narrow_band_least_squares_b200.engine import Engine  # noqa: E402
from narrow_band_least_squares_b200.synth import plane_wave_record  # noqa: E402
from oracle import nbls_oracle as orc  # noqa: E402
from test_gpu_wide import adversarial_record  # noqa: E402

eng = Engine(0)
out = {"band": "XF_TAU32 = 2e-5 (half-width 1e-5)", "plane_wave_c3": {}, "adversarial": {}}


def audited(filt, n, hop, nwin, two_stage=1, method=_capi.XCORR_AUTO):
    eng.set_option(_capi.OPT_SCREEN_AUDIT, 1)
    eng.set_option(_capi.OPT_TWO_STAGE, two_stage)
    eng.xcorr_audit(reset=True)
    eng.xcorr_stats(reset=True)
    try:
        eng.xcorr(filt, [n], [hop], [nwin], method=method)
    finally:
        eng.set_option(_capi.OPT_SCREEN_AUDIT, 0)
        eng.set_option(_capi.OPT_TWO_STAGE, 1)
    err, nl = eng.xcorr_audit(reset=True)
    st = eng.xcorr_stats(reset=True)
    return {"max_err_all_lags": err, "lags_compared": nl, "max_err_at_maximum": st["max_screen_err"],
            "candidates_per_pair_window": st["verified"] / max(1, nwin * filt.shape[1] * (filt.shape[1] - 1) // 2),
            "time_domain_items": st["time_domain"]}


j = config_job("c3")
rec = plane_wave_record(1003, 25, 120000, 100.0, 1.0)
for b in (0, 5, 10, 15, 19):
    sos = [orc.design_sos("butter", j["freqlist"][b], j["freqlist"][b + 1], rec.fs, 2)]
    filt = eng.filter_bands(rec.data, rec.fs, sos)
    n = int(j["winlen_list"][b] * rec.fs)
    hop = n // 2
    nwin = len(np.arange(0, rec.npts - n, hop))
    for ts in (1, 0):
        out["plane_wave_c3"][f"band{b}_N{n}_{'two_stage' if ts else 'single_stage'}"] = audited(filt, n, hop, nwin, ts)
for kind in ("dc_tone", "spike", "ramp", "sines", "noise"):
    for n in (1500, 3000, 6000):
        hop = n // 2
        nwin = 120
        x = adversarial_record(kind, 77, 16, hop * (nwin + 1) + 1, 100.0, n)
        out["adversarial"][f"{kind}_N{n}"] = audited(x[None], n, hop, nwin, 0, _capi.XCORR_FFT32)
print(json.dumps(out))
