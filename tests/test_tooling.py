"""CPU tests of the measurement tooling: the ncu launch-list parser behind `roofline.traffic`, the long-slice parity
counter's CPU side against the oracle, the block-wise record generator, and the staged-reference archive."""
import importlib.util
import json
import os
import subprocess
import sys

import numpy as np

from oracle import nbls_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_ncu_traffic_parser_on_the_committed_launch_list(tmp_path):
    """scripts/ncu_traffic.py turns the committed ncu launch list into the per-group DRAM bytes bench.py reads."""
    csv = os.path.join(ROOT, "profiles", "r2_ncu_launch_list_c3_2h.csv")
    env = dict(os.environ)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_traffic.py"), csv, "--key", "test_key",
                          "--scale", "12", "--summary", str(tmp_path / "s.csv")], capture_output=True, text=True, env=env)
    try:
        assert out.returncode == 0, out.stderr
        got = json.loads(out.stderr.strip().splitlines()[-1])
        # spectra are written and re-read, the pair kernel reads them through L2; the timed path asks for no per-pair cmax,
        # so the dense verify kernel does not run and the item kernel + median selection re-read few samples
        assert 5e10 < got["xf_pair"] < 7e10 and 7e10 < got["xf_spectra"] < 1e11 and 5e9 < got["xf_verify_items"] < 4e10
        assert "xf_verify" not in got
        assert {"filter_fwd_phase1", "filter_fwd_phase3", "filter_bwd_phase1", "filter_bwd_phase3"} <= set(got)
        rows = open(tmp_path / "s.csv").read().splitlines()
        assert rows[0].startswith("group,launches") and rows[1].startswith("xf_pair,")
    finally:
        p = os.path.join(ROOT, "profiles", "ncu_traffic.json")        # the script merged a test key: take it out again
        d = json.load(open(p))
        d.pop("test_key", None); d.pop("test_key__meta", None)
        json.dump(d, open(p, "w"), indent=1, sort_keys=True)


def test_bench_reads_traffic_for_the_headline_key():
    d = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    key = d["c3_24h_auto"]
    assert key["xf_pair"] > 0 and key["filter_bwd_phase1"] < key["filter_bwd_phase3"]
    # the filter groups come from the full-length pass (not scaled): ~1.17x the algorithmic 105.4 GB in all
    filt = sum(v for k, v in key.items() if k.startswith("filter_"))
    assert 1.05e11 < filt < 1.35e11


def test_parity_counter_fft_screen_equals_oracle_argmax():
    """scripts/parity_count.py screens with a float64 FFT correlation: on ordinary data its arg-max is np.correlate's."""
    pc = _load(os.path.join(ROOT, "scripts", "parity_count.py"), "parity_count")
    from narrow_band_least_squares_b200.synth import plane_wave_record
    rec = plane_wave_record(3, 5, 6000, 50.0, 0.4, dead_channel=2)
    filt, _ = orc.filter_record(rec.data, 50.0, "butter", 0.5, 1.0, 2)
    n, hop, starts = orc.window_grid(rec.npts, 50.0, 20.0, 0.5)
    pairs, _, _, _ = orc.co_array(rec.rij)
    lag, *_ = orc.correlate_windows(np.ascontiguousarray(filt.T), starts, n, 50.0, pairs)
    got = pc.band_lags_fft(filt, n, hop, len(starts), pairs)
    assert np.array_equal(got.T, lag)
    assert pc.exact_lag(filt[0, :n].copy(), filt[1, :n].copy()) == lag[0, 0]


def test_blocked_record_segments_agree_with_the_whole():
    """Any rank can synthesise any segment of a block-wise record (synth.blocked_record): segments equal the whole."""
    from narrow_band_least_squares_b200.synth import blocked_record
    whole = blocked_record(11, 4, 0, 9000, 100.0, 0.3, block=4000, flip_channel=1, shift_channel=2)
    for a, b in ((0, 4000), (2500, 8100), (3999, 4001), (8000, 9000)):
        part = blocked_record(11, 4, a, b, 100.0, 0.3, block=4000, flip_channel=1, shift_channel=2)
        assert np.array_equal(part.data, whole.data[:, a:b])
        assert np.array_equal(part.rij, whole.rij)
    quiet = blocked_record(11, 4, 0, 9000, 100.0, 0.3, block=4000, noise_every_min=1)      # every minute noise-only
    assert abs(quiet.data.std() - 0.3) < 0.02


def test_staged_reference_archive_runs_the_reference_driver():
    """oracle/refrun.stage_reference packs the hot-path files; the reference's serial driver runs from the archive and
    agrees with the oracle (build container only: needs /root/reference to pack from)."""
    import pytest
    from oracle import refrun
    if not os.path.isdir("/root/reference"):
        pytest.skip("no /root/reference here (the archive arrives with the snapshot)")
    assert refrun.stage_reference() == len(refrun.HOT_PATH_FILES)
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r)\n"
        "from oracle import refrun, nbls_oracle as orc\n"
        "from narrow_band_least_squares_b200.synth import plane_wave_record\n"
        "import contextlib, io\n"
        "ref = refrun.import_reference(refrun.ARCHIVE)\n"
        "assert '.zip' in ref[1].__file__\n"
        "rec = plane_wave_record(5, 4, 2400, 40.0, 0.2)\n"
        "fl = orc.get_freqlist_log(0.5, 5.0, 2)\n"
        "with contextlib.redirect_stdout(io.StringIO()):\n"
        "    out = refrun.run_reference_driver(ref, rec, fl, [10.0, 10.0], 0.5, 1.0, 2, parallel=False)\n"
        "tvec = np.arange(rec.npts) / rec.fs / 86400.0\n"
        "o = orc.nbls_oracle(rec.data, rec.fs, tvec, rec.rij, fl, [10.0, 10.0], 0.5, 1.0)\n"
        "for b in range(2):\n"
        "    w = out[6][b]\n"
        "    assert w == o['num_compute'][b]\n"
        "    assert np.allclose(out[0][b, :w], o['vel'][b, :w], rtol=1e-12, equal_nan=True)\n"
        "print('ok')\n" % ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stderr[-2000:]


def test_committed_headline_line_keeps_the_bench_contract():
    """The last headline bench line (profiles/r2_bench_c3_final4.json, unedited stdout of bench.py) carries every key the
    bench contract names, with the metric and workload of BASELINE.json."""
    d = json.loads(open(os.path.join(ROOT, "profiles", "r2_bench_c3_final4.json")).read())
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks", "parity"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["warmup"] >= 3 and d["higher_is_better"] is True and d["gpu_launches"] > 0
    assert d["unit"] == base.get("unit", d["unit"]) and "workload" in d["config"] and d["config"]["workload"].startswith("c3")
    r = d["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(r) and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    c = d["cpu_baseline"]
    assert {"value", "unit", "cores", "kind", "sample"} <= set(c) and c["kind"] == "reference" and c["cores"] >= 1
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and 0 < e["value"] <= 1.05 * d["value"]
    assert abs(d["value"] - d["config"]["estimates_per_step"] / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-6
    assert not ({"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(d["clocks"]["reasons"]))
    p = d["parity"]
    assert p["vs_oracle"]["lag_mismatches"] == 0 and p["timed_path_vs_cmax_path"]["lag_mismatches"] == 0
    assert p["timed_path_vs_cmax_path"]["estimates_bit_identical"] and p["timed_path_vs_cmax_path"]["max_abs_dmdccm"] <= 1e-15
