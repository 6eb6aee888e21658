"""Pin the CPU oracle (oracle/nbls_oracle.py) against outputs of the UNMODIFIED reference.

The fixtures under tests/golden/ were produced in the build container by
oracle/run_reference.py importing /root/reference (lts_array.ltsva, helpers.filter_data,
narrow_band_least_squares).  The reference itself has no tests or golden vectors.
"""
import hashlib
import os

import numpy as np
import pytest

from narrow_band_least_squares_b200.synth import plane_wave_record
from oracle import nbls_oracle as orc
from oracle.run_reference import CASES

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _load(name):
    return np.load(os.path.join(GOLD, name + ".npz"), allow_pickle=False)


def test_lcg_subset_stream_matches_reference():
    g = _load("tables")
    for p in (6, 10, 15, 28, 66, 300):
        sub, seeds = orc.lcg_subsets(p, 520)
        assert np.array_equal(sub, g[f"randset_{p}"])
        assert np.array_equal(seeds, g[f"randseed_{p}"])
    # SURVEY A.6-5 known answers
    assert orc.lcg_subsets(15, 6)[0].tolist() == [[0, 12], [3, 13], [9, 1], [3, 8], [0, 4], [1, 2]]
    assert orc.lcg_subsets(6, 6)[1][-1] == 20411


def test_correction_factors_match_reference():
    g = _load("tables")
    for p, alpha, h, raw, rew1 in g["factors"]:
        p = int(p)
        assert orc.h_value(p, alpha) == int(h)
        assert orc.raw_corfactor(2, p, alpha) * orc.raw_consfactor(int(h), p) == raw
        assert orc.rew_corfactor(2, p, alpha) == rew1
    for p in (6, 10, 15, 28, 66, 300):
        tab = g[f"rewcons_{p}"]
        mine = np.array([np.nan] + [orc.rew_consfactor(m, p) for m in range(1, p)] + [np.nan])
        assert np.array_equal(tab[1:p], mine[1:p], equal_nan=True)


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_stage_by_stage(name):
    g = _load(name)
    skw, pkw = CASES[name]
    rec = plane_wave_record(**skw)
    assert _sha(rec.data) == str(g["input_sha"]), "synthetic generator drifted from the fixture inputs"
    assert np.allclose(rec.rij, g["b0_rij"], rtol=0, atol=1e-8)      # Vincenty: unpinned (obspy absent)
    fs = rec.fs
    tvec = (rec.starttime + np.arange(rec.npts) / fs) / 86400.0
    fl, wl = g["freqlist"], g["winlen_list"]
    for b in range(int(g["nbands"])):
        filt, sos = orc.filter_record(rec.data, fs, "butter", fl[b], fl[b + 1], int(g["order"]))
        assert np.array_equal(sos, g[f"b{b}_sos"])
        assert _sha(filt) == str(g[f"b{b}_filt_sha"])          # same scipy calls as the stub: bit-identical
        res = orc.ltsva_oracle(filt, fs, tvec, g["b0_rij"], float(wl[b]), float(g["winover"]), float(g["alpha"]))
        assert res["n"] == int(g[f"b{b}_winlensamp"]) and res["hop"] == int(g[f"b{b}_sampinc"])
        assert np.array_equal(res["starts"], g[f"b{b}_intervals"])
        assert np.array_equal(res["lag"], g[f"b{b}_lag"])                      # integer lags: exact
        assert np.array_equal(res["tau"], g[f"b{b}_tau"])
        assert np.array_equal(res["cmax"], g[f"b{b}_cmax"], equal_nan=True)
        assert np.array_equal(res["mdccm"], g[f"b{b}_mdccm"], equal_nan=True)
        assert np.array_equal(res["mad"], g[f"b{b}_mad"])
        assert np.array_equal(res["t"], g[f"b{b}_t"])
        for k in ("vel", "baz", "sigma_tau"):
            assert np.array_equal(np.isnan(res[k]), np.isnan(g[f"b{b}_{k}"])), k
        assert np.allclose(res["vel"], g[f"b{b}_vel"], rtol=1e-9, atol=0, equal_nan=True)
        assert np.allclose(res["baz"], g[f"b{b}_baz"], rtol=0, atol=1e-7, equal_nan=True)
        assert np.allclose(res["sigma_tau"], g[f"b{b}_sigma_tau"], rtol=1e-7, atol=1e-12, equal_nan=True)
        if float(g["alpha"]) != 1.0:
            assert res["h"] == int(g[f"b{b}_h"])
            assert np.array_equal(res["element_weights"], g[f"b{b}_element_weights"])   # LTS flags: exact
            assert np.allclose(res["slowness_coeffs"], g[f"b{b}_slowness_coeffs"], rtol=1e-9, atol=1e-14,
                               equal_nan=True)
            keys = [k for k in res["stdict"] if k != "size"]
            assert keys == [str(k) for k in g[f"b{b}_stdict_keys"]]
            off = g[f"b{b}_stdict_off"]
            for i, k in enumerate(keys):
                assert np.array_equal(res["stdict"][k], g[f"b{b}_stdict_vals"][off[i]:off[i + 1]])
            assert res["stdict"]["size"] == int(g[f"b{b}_stdict_size"])


@pytest.mark.parametrize("name", ["g1_ols_c4", "g6_lts_c6_flip"])
def test_oracle_driver_matches_reference_driver(name):
    g = _load(name)
    skw, pkw = CASES[name]
    rec = plane_wave_record(**skw)
    tvec = (rec.starttime + np.arange(rec.npts) / rec.fs) / 86400.0
    out = orc.nbls_oracle(rec.data, rec.fs, tvec, g["b0_rij"], g["freqlist"], list(g["winlen_list"]),
                          float(g["winover"]), float(g["alpha"]), order=int(g["order"]))
    assert out["num_compute"] == g["nbls_num_compute"].tolist()
    assert out["vel"].shape == g["nbls_vel"].shape             # the vector_len sizing quirk (SURVEY a-14)
    assert np.allclose(out["vel"], g["nbls_vel"], rtol=1e-9, equal_nan=True)
    assert np.allclose(out["baz"], g["nbls_baz"], rtol=0, atol=1e-7, equal_nan=True)
    assert np.array_equal(out["mdccm"], g["nbls_mdccm"], equal_nan=True)
    assert np.array_equal(out["t"], g["nbls_t"], equal_nan=True)
    if float(g["alpha"]) == 1.0:
        assert np.allclose(out["sigma_tau"], g["nbls_sig_tau"], rtol=1e-7, atol=1e-12, equal_nan=True)
        assert out["stdict_all"] is None
    else:
        keys = [k for k in out["stdict_all"] if k != "size"]
        assert keys == [str(k) for k in g["nbls_stdict_keys"]]
