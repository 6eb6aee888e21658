"""CPU tests (no GPU): the C-ABI library loads and exports every symbol include/nbls.h declares, and the host-side
logic (window grid, co-array, LTS tables, result assembly, time-segment sharding) follows the reference."""
import os
import re

import numpy as np
import pytest

from narrow_band_least_squares_b200 import _capi, helpers, lts_tables, sharding
from narrow_band_least_squares_b200.engine import window_grid
from narrow_band_least_squares_b200.geometry import CoArray, get_rij
from narrow_band_least_squares_b200.ltsva import build_stdict
from narrow_band_least_squares_b200.narrow_band_least_squares import reference_vector_len
from oracle import nbls_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def test_capi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "nbls.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(nbls_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = _capi.load()                      # raises if the library was not built
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} is declared in nbls.h but not exported by libnbls.so"
    assert declared == set(_capi.EXPORTS), "ctypes prototypes and nbls.h disagree"
    assert lib.nbls_version() >= 100


def test_capi_option_numbers_match_the_header():
    """NBLS_OPT_* of include/nbls.h and _capi.OPT_* (what Engine.set_option and bench.py --opt use) are the same numbers,
    every one below NBLS_OPT_COUNT; NBLS_XCORR_STATIC_MAX_ENTRIES is stated in the header."""
    hdr = open(os.path.join(ROOT, "include", "nbls.h")).read()
    defs = {k: int(v) for k, v in re.findall(r"#define\s+NBLS_OPT_([A-Z0-9_]+)\s+(\d+)", hdr)}
    count = defs.pop("COUNT")
    assert len(defs) >= 11 and sorted(defs.values()) == list(range(len(defs))) and max(defs.values()) < count
    for name, value in defs.items():
        assert getattr(_capi, "OPT_" + name) == value, name
    assert re.search(r"#define\s+NBLS_XCORR_STATIC_MAX_ENTRIES\s+262144", hdr)


def test_no_cuda_device_fails_loudly():
    import ctypes as C
    lib = _capi.load()
    n = C.c_int32(-1)
    assert lib.nbls_device_count(C.byref(n)) == 0
    if n.value == 0:                        # the build container: no CPU fallback may exist
        ctx = C.c_void_p()
        assert lib.nbls_create(C.byref(ctx), 0) != 0
        assert b"no CPU fallback" in lib.nbls_last_error()
        from narrow_band_least_squares_b200.engine import Engine
        with pytest.raises(_capi.NBLSError):
            Engine(0)


def test_fft_length_table():
    lib = _capi.load()
    assert lib.nbls_xcorr_fft_length(3000, -1) == 6144        # N + (N-1) = 5999 -> 6 * 1024
    assert lib.nbls_xcorr_fft_length(6000, -1) == 12288
    assert lib.nbls_xcorr_fft_length(1200, -1) == 2560        # 2399 -> 5 * 512
    assert lib.nbls_xcorr_fft_length(5000, -1) == 10240       # 9999 -> 5 * 2048
    assert lib.nbls_xcorr_fft_length(3000, 200) == 4096       # restricted lag: N + 200 points suffice
    assert lib.nbls_xcorr_fft_length(6145, -1) == 0           # beyond one SM's shared memory: time-domain kernel
    assert lib.nbls_xcorr_fft_length(0, -1) == 0


@pytest.mark.parametrize("npts,fs,wl,ov", [(144000, 40.0, 30.0, 0.5), (8640000, 100.0, 30.0, 0.5), (6000, 50.0, 20.0, 0.75),
                                            (3000, 100.0, 30.0, 0.5), (2999, 100.0, 30.0, 0.5)])
def test_window_grid_matches_databin(npts, fs, wl, ov):
    n, hop, w = window_grid(npts, fs, wl, ov)
    n2, hop2, starts = orc.window_grid(npts, fs, wl, ov)      # lts_data_class.py:58-64
    assert (n, hop, w) == (n2, hop2, len(starts))
    if (npts, fs, wl) == (144000, 40.0, 30.0):
        assert w == 238 and reference_vector_len(npts, fs, [wl], ov) == 239
    if (npts, fs, wl) == (8640000, 100.0, 30.0):
        assert w == 5758


def test_helpers_follow_reference_lists():
    fl, nb, _ = helpers.get_freqlist(0.5, 5.0, "log", 8)
    assert nb == 8 and np.allclose(fl, orc.get_freqlist_log(0.5, 5.0, 8))
    assert helpers.get_winlenlist("constant", 4, 30, 60, 30) == [30, 30, 30, 30]
    wl = helpers.get_winlenlist("adaptive", 20, 30, 60, 30)
    assert len(wl) == 20 and wl[0] == 60 and wl[-1] == 30
    sos = helpers.design_sos("butter", 0.5, 0.8, 100.0, 2)
    assert np.array_equal(sos, orc.design_sos("butter", 0.5, 0.8, 100.0, 2))
    wlen, up, down = helpers.taper_ramps(144000, 0.01)
    wlen2, up2, down2 = orc.taper_weights(144000, 0.01)
    assert wlen == wlen2 == 1440 and np.array_equal(up, up2) and np.array_equal(down, down2)


def test_coarray_and_lts_tables_match_golden():
    g = np.load(os.path.join(GOLD, "tables.npz"))
    for p in (6, 10, 15, 28, 66, 300):
        assert np.array_equal(lts_tables.subset_stream(p, 520), g[f"randset_{p}"])
    for p, alpha, h, raw, rew1 in g["factors"]:
        hh, r, w1, tab = lts_tables.lts_constants(int(p), float(alpha))
        assert hh == int(h) and r == raw and w1 == rew1
        assert np.array_equal(tab[1:int(p)], g[f"rewcons_{int(p)}"][1:int(p)], equal_nan=True)
    rng = np.random.default_rng(3)
    rij = rng.standard_normal((2, 7))
    co = CoArray(rij)
    pairs, xij, mad, xstd = orc.co_array(rij)
    assert np.array_equal(co.idx, pairs) and np.array_equal(co.xij, xij)
    assert np.array_equal(co.xij_mad, mad) and np.array_equal(co.xij_standardized, xstd)
    with pytest.raises(RuntimeError):
        CoArray(np.array([[0.0, 1.0, 2.0], [0.0, 1.0, 2.0]])).check_rank()


def test_stdict_assembly_matches_reference_rule():
    w = np.ones((6, 3), dtype=np.uint8)
    w[1, 0] = 0
    w[:, 2] = 0                                   # skipped window: every pair flagged
    t = np.array([1.0, 2.0, 3.0])
    idx = orc.co_array(np.random.default_rng(0).standard_normal((2, 4)))[0]
    sd = build_stdict(w, t, np.asarray(idx), 4)
    ref = orc.build_stdict(w, t, idx, 4)
    assert list(sd) == list(ref) and sd["size"] == 4
    for k in sd:
        assert np.array_equal(sd[k], ref[k])


def test_vincenty_against_oracle_stub():
    lat = [64.0, 64.001, 63.9995, 64.0007]
    lon = [-147.0, -147.002, -146.9991, -147.0012]
    rij = get_rij(lat, lon)
    assert rij.shape == (2, 4) and np.allclose(rij.mean(axis=1), 0.0, atol=1e-12)
    d01 = np.hypot(*(rij[:, 0] - rij[:, 1]))
    assert 0.12 < d01 < 0.16                      # ~111 m north, ~97 m west


# ---- time-segment sharding --------------------------------------------------------------------
def _bands():
    fs = 100.0
    fl = orc.get_freqlist_log(0.5, 5.0, 4)
    sos = [orc.design_sos("butter", fl[b], fl[b + 1], fs, 2) for b in range(4)]
    return fs, fl, sos


def test_shards_partition_windows_and_keep_halos():
    fs, fl, sos = _bands()
    npts = 600000
    grids = [window_grid(npts, fs, wl, 0.5) for wl in (60.0, 50.0, 40.0, 30.0)]
    n = [g[0] for g in grids]; hop = [g[1] for g in grids]; cnt = [g[2] for g in grids]
    halo = sharding.filter_halo(sos)
    assert halo == max(int(np.ceil(80 * np.log(2) / -np.log(sharding.pole_radius(s)))) for s in sos)
    for world in (1, 2, 3, 8):
        shards = [sharding.Shard(npts, n, hop, cnt, r, world, halo) for r in range(world)]
        for b in range(4):
            lo = [int(s.w_lo[b]) for s in shards]; hi = [int(s.w_hi[b]) for s in shards]
            assert lo[0] == 0 and hi[-1] == cnt[b]
            assert all(hi[r] == lo[r + 1] for r in range(world - 1))          # disjoint, complete, ordered
        for s in shards:
            assert s.seg_lo % 16 == 0 and 0 <= s.seg_lo < s.seg_hi <= npts
            for b in range(4):
                if s.win_count[b]:
                    first = int(s.win_first[b]) + s.seg_lo
                    last_end = first + (int(s.win_count[b]) - 1) * hop[b] + n[b]
                    assert first == int(s.w_lo[b]) * hop[b]
                    assert first - s.seg_lo >= (halo if s.rank > 0 else 0) - 15
                    assert s.seg_hi - last_end >= (halo if s.rank < world - 1 else 0) or s.seg_hi == npts
        assert shards[0].seg_lo == 0 and shards[-1].seg_hi == npts


def test_halo_makes_segment_filter_equal_whole_record_filter():
    """The oracle filter on a halo-padded segment agrees with the whole-record filter inside the owned windows
    to below FP64 rounding, which is what makes sharded lags equal unsharded ones."""
    fs, fl, sos = _bands()
    rng = np.random.default_rng(11)
    npts = 400000
    x = rng.standard_normal((2, npts))
    halo = sharding.filter_halo(sos)
    whole = [orc.filter_record(x, fs, "butter", fl[b], fl[b + 1], 2, taper=None)[0] for b in range(4)]
    grids = [window_grid(npts, fs, 30.0, 0.5)] * 4
    sh = sharding.Shard(npts, [g[0] for g in grids], [g[1] for g in grids], [g[2] for g in grids], 1, 3, halo)
    seg = x[:, sh.seg_lo:sh.seg_hi]
    for b in range(4):
        part = orc.filter_record(seg, fs, "butter", fl[b], fl[b + 1], 2, taper=None)[0]
        a = int(sh.win_first[b]); z = a + (int(sh.win_count[b]) - 1) * grids[b][1] + grids[b][0]
        ref = whole[b][:, sh.seg_lo + a:sh.seg_lo + z]
        assert np.max(np.abs(part[:, a:z] - ref)) <= 1e-13 * np.max(np.abs(ref))


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        counts = [3 + rank, 2, 0 if rank == 0 else 4]                 # windows per band on this rank
        e = sum(counts)
        base = 1000 * rank
        local = {"vel": base + np.arange(e, dtype=np.float64), "lag": (base + np.arange(e * 5, dtype=np.int32)).reshape(e, 5)}
        out = sharding.gather_results(local, counts)
        q.put((rank, {k: v.tolist() for k, v in out.items()}))
    finally:
        dist.destroy_process_group()


def test_gather_results_world2_gloo():
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # global band-major order: band 0 of rank 0, band 0 of rank 1, band 1 of rank 0, ...
    r0 = np.arange(5, dtype=float); r1 = 1000 + np.arange(10, dtype=float)
    exp = np.concatenate([r0[0:3], r1[0:4], r0[3:5], r1[4:6], r1[6:10]])
    for rank in (0, 1):
        assert np.array_equal(np.array(got[rank]["vel"]), exp)
        assert np.array(got[rank]["lag"]).shape == (15, 5)


def _gloo_packed_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        counts = [[3, 2, 0], [4, 2, 4]]                              # windows per band on rank 0 / rank 1
        pg = sharding.PackedGather(counts, rank, npairs=5, want_lag=True, want_weights=True, device="cpu")
        e, base = pg.e_local, 1000 * rank
        o = pg.outputs()
        assert o.vel == pg.local.data_ptr() + 2 * 8 * pg.e_max and o.slowness == pg.local.data_ptr() + 5 * 8 * pg.e_max
        # what the kernels would write through those pointers: column arrays of e_local rows, slowness [e][2] interleaved
        pg.local[2, :e] = base + torch.arange(e, dtype=torch.float64)
        pg.local[5:7].reshape(-1)[:2 * e] = (base + torch.arange(e, dtype=torch.float64)).repeat_interleave(2) + \
            torch.tensor([0.25, 0.5], dtype=torch.float64).repeat(e)
        pg.local_lag[:e] = (base + torch.arange(e * 5, dtype=torch.int32)).reshape(e, 5)
        pg.local_w[:e] = rank + 1
        out = pg.gather()
        q.put((rank, {k: v.tolist() for k, v in out.items()}))
    finally:
        dist.destroy_process_group()


def test_packed_gather_world2_gloo():
    """One all_gather_into_tensor of the packed estimate block + a precomputed permutation = global band-major order."""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_packed_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    r0 = np.arange(5, dtype=float); r1 = 1000 + np.arange(10, dtype=float)
    exp = np.concatenate([r0[0:3], r1[0:4], r0[3:5], r1[4:6], r1[6:10]])
    for rank in (0, 1):
        assert np.array_equal(np.array(got[rank]["vel"]), exp)
        sl = np.array(got[rank]["slowness"])
        assert sl.shape == (15, 2) and np.array_equal(sl[:, 0], exp + 0.25) and np.array_equal(sl[:, 1], exp + 0.5)
        lag = np.array(got[rank]["lag"])
        assert lag.shape == (15, 5) and np.array_equal(lag[:, 0], 5 * exp - 4000 * (exp >= 1000))
        assert np.array_equal(np.array(got[rank]["weights"])[:, 0], 1 + (exp >= 1000))


@pytest.mark.parametrize("name", ["g1_ols_c4", "g2_lts_c6", "g4_lts_c12", "g5_ols_c5_overlap75", "g6_lts_c6_flip"])
def test_slowness_ellipse_uncertainties_match_reference(name):
    _ellipse_check(name)


def _ellipse_check(name):
    """conf_int_vel / conf_int_baz (uncertainty.py) against the values the unmodified reference produced
    (lts_classes.py:715-759, :938-968) for EVERY window — also those whose confidence ellipse contains the origin,
    where the reference keeps the real parts of complex tangency roots (rthEllipse :617-633)."""
    import os
    from narrow_band_least_squares_b200 import uncertainty
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", name + ".npz"), allow_pickle=True)
    alpha, checked, inside = float(g["alpha"]), 0, 0
    seen = []
    orig = uncertainty._tangent_half_angle_inside

    def spy(a, b, x0, y0):
        seen.append(len(a))
        return orig(a, b, x0, y0)
    uncertainty._tangent_half_angle_inside = spy
    try:
        for b in range(int(g["nbands"])):
            vel, baz, sig = g[f"b{b}_vel"], g[f"b{b}_baz"], g[f"b{b}_sigma_tau"]
            s = 1.0 / vel
            slow = np.stack([s * np.sin(np.radians(baz)), s * np.cos(np.radians(baz))], axis=1)
            w = g[f"b{b}_element_weights"].T if alpha != 1.0 else None
            civ, cib = uncertainty.slowness_ellipse_ci(slow, sig, g["b0_xij"], w)
            rv, rb = g[f"b{b}_conf_int_vel"], g[f"b{b}_conf_int_baz"]
            assert np.array_equal(np.isnan(civ), np.isnan(rv)) and np.array_equal(np.isnan(cib), np.isnan(rb))
            m = np.isfinite(rv)
            assert np.allclose(civ[m], rv[m], rtol=1e-5, atol=0)
            assert np.allclose(cib[m], rb[m], rtol=0, atol=1e-6)
            checked += int(m.sum())
    finally:
        uncertainty._tangent_half_angle_inside = orig
    inside = sum(seen)
    assert checked > 10
    return inside


def test_slowness_ellipse_inside_origin_windows_are_covered():
    """The fixtures do contain windows whose ellipse holds the origin (noise-only minutes), so the test above pins them."""
    total = sum(_ellipse_check(n) for n in
                ["g1_ols_c4", "g2_lts_c6", "g4_lts_c12", "g5_ols_c5_overlap75", "g6_lts_c6_flip"])
    assert total > 0


def test_slowness_ellipse_known_geometry():
    """Circle of radius r centred at distance d on the x axis: distances d -+ r, tangent half-angle asin(r/d)."""
    from narrow_band_least_squares_b200.uncertainty import slowness_ellipse_ci
    xij = np.array([[1.0, 0.0], [0.0, 1.0], [-1.0, 0.0], [0.0, -1.0]])      # X^T X = 2 I
    chi = np.sqrt(-2.0 * np.log(0.1))
    sig = np.array([0.1, 0.2])
    r = chi * sig / np.sqrt(2.0)
    d = 3.0
    civ, cib = slowness_ellipse_ci(np.array([[d, 0.0], [0.0, d]]), sig, xij)
    assert np.allclose(civ, 0.5 * np.abs(1.0 / (d - r) - 1.0 / (d + r)), rtol=1e-12)
    assert np.allclose(cib, np.degrees(np.arcsin(r / d)), rtol=1e-10)


def _helpers_golden():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "helpers.npz"), allow_pickle=True)


def test_freqlist_every_band_layout_matches_reference():
    """get_freqlist for 'linear', 'log', 'octave', '2_octave_over', 'onethird_octave', 'octave_linear' against the
    lists the unmodified reference produced (helpers.py:8-79; fixture written by oracle/run_reference.py)."""
    g = _helpers_golden()
    n = 0
    for kind in ("linear", "log", "octave", "2_octave_over", "onethird_octave", "octave_linear"):
        for ci in range(3):
            fmin, fmax, nb, nbc, fmc = g[f"freq_{kind}_{ci}_meta"]
            if nbc < 0:      # the reference divides by zero (octave_linear without a linear part)
                with pytest.raises(ZeroDivisionError):
                    helpers.get_freqlist(float(fmin), float(fmax), kind, int(nb))
                continue
            fl, nbands_calc, fmax_calc = helpers.get_freqlist(float(fmin), float(fmax), kind, int(nb))
            assert np.array_equal(np.asarray(fl, dtype=np.float64), g[f"freq_{kind}_{ci}"]), (kind, ci)
            assert nbands_calc == int(nbc) and fmax_calc == fmc
            n += 1
    assert n >= 15


def test_winlenlist_matches_reference():
    g = _helpers_golden()
    for ci, kind in enumerate(("constant", "adaptive", "adaptive")):
        nb, wl, w1, wx = g[f"winlen_{ci}_meta"]
        got = helpers.get_winlenlist(kind, int(nb), wl, w1, wx)
        assert np.array_equal(np.asarray(got, dtype=np.float64), g[f"winlen_{ci}"])


def test_results_txtfile_is_byte_identical_to_reference(tmp_path):
    """write_txtfile produces the reference's file byte for byte (helpers.py:161-181); read_txtfile returns what
    the reference's reader returns (helpers.py:184-235)."""
    g = _helpers_golden()
    d = str(tmp_path) + "/"
    helpers.write_txtfile(d, "mine", g["txt_vel"], g["txt_baz"], g["txt_mdccm"], g["txt_t"], g["txt_freqlist"],
                          [int(v) for v in g["txt_ncl"]])
    assert open(d + "mine.txt", "rb").read() == g["txt_bytes"].tobytes()
    back = helpers.read_txtfile(d, "mine")
    for k, v in zip(("vel", "baz", "mdccm", "t", "freqlist", "ncl"), back[:6]):
        assert np.array_equal(np.asarray(v, dtype=np.float64), g[f"txt_back_{k}"], equal_nan=True), k
    assert np.array_equal(np.array(back[6:9], dtype=np.float64), g["txt_back_scalars"])


def test_remove_elements_edits_the_callers_lists_like_the_reference():
    """lts_data_class.py:34-40: the listed elements (python numbering, any order) leave the caller's stream and coordinate
    lists in place; the data matrix, element names and co-array are built from what remains."""
    from narrow_band_least_squares_b200.ltsva import DataBin
    from narrow_band_least_squares_b200.stream import stream_from_array
    rng = np.random.default_rng(3)
    x = rng.standard_normal((6, 4000))
    st = stream_from_array(x, 100.0, stations=[f"I53H{k + 1}" for k in range(6)])
    lat = [64.0 + 1e-3 * k for k in range(6)]
    lon = [-147.0 - 2e-3 * k * k for k in range(6)]
    lat_all, lon_all = list(lat), list(lon)
    d = DataBin(10.0, 0.5, 1.0)
    d.build_data_arrays(st, lat, lon, remove_elements=[4, 1])
    keep = [0, 2, 3, 5]
    assert len(st) == 4 and lat == [lat_all[k] for k in keep] and lon == [lon_all[k] for k in keep]
    assert d.nchans == 4 and d.element_names == [f"{k + 1:02d}" for k in keep]      # Trace.stats.location
    assert np.array_equal(d.data_ct, x[keep])
    ref = DataBin(10.0, 0.5, 1.0)
    ref.build_data_arrays(stream_from_array(x[keep], 100.0, stations=[f"I53H{k + 1}" for k in keep]),
                          [lat_all[k] for k in keep], [lon_all[k] for k in keep])
    assert np.array_equal(d.rij, ref.rij) and d.nits == ref.nits == (4000 - 1000 + 499) // 500
    with pytest.raises(RuntimeError):                    # fewer than 3 elements left (lts_data_class.py:76-80)
        DataBin(10.0, 0.5, 1.0).build_data_arrays(stream_from_array(x[:4], 100.0), lat_all[:4], lon_all[:4],
                                                   remove_elements=[0, 1])



def test_deal_arrays_by_array_then_segment():
    """Config 5 sharding rule: arrays first, time segments only when ranks outnumber arrays."""
    from narrow_band_least_squares_b200.batch import deal_arrays
    items = deal_arrays(64, 8)
    assert all(len(it) == 8 and all(s == 0 and g == 1 for _, s, g in it) for it in items)
    assert sorted(a for it in items for a, _, _ in it) == list(range(64))
    assert [a for a, _, _ in items[3]] == list(range(3, 64, 8))
    items = deal_arrays(3, 8)                     # 8 ranks, 3 arrays: groups of 3, 3 and 2 ranks share an array by segment
    assert [it[0] for it in items] == [(0, 0, 3), (0, 1, 3), (0, 2, 3), (1, 0, 3), (1, 1, 3), (1, 2, 3), (2, 0, 2), (2, 1, 2)]
    assert deal_arrays(5, 1) == [[(a, 0, 1) for a in range(5)]]


def test_lazy_verification_median_rule():
    """The rule of xf_median_select_kernel (csrc/xcorr_fft.cu), restated in numpy: stored values v within eps of the exact
    ones e; after the approximate entries within 2 eps of either middle order statistic of v are replaced by their exact
    values, np.nanmedian of the stored array equals np.nanmedian(e) - whatever the errors were (worst cases included:
    values clustered inside the band, errors at +-eps, NaNs, entries that are exact already)."""
    rng = np.random.default_rng(5)
    eps = 1e-5
    for trial in range(4000):
        p = int(rng.choice([6, 15, 66, 299, 300]))
        kind = trial % 4
        if kind == 0:
            e = rng.uniform(0.0, 1.0, p)
        elif kind == 1:                               # everything inside a few bands of the median
            e = 0.7 + rng.uniform(-4 * eps, 4 * eps, p)
        elif kind == 2:                               # ties and near-ties
            e = np.round(rng.uniform(0.5, 0.5 + 20 * eps, p) / (eps / 2)) * (eps / 2)
        else:
            e = np.concatenate([rng.uniform(0, 1, p // 2), 0.5 + rng.uniform(-2 * eps, 2 * eps, p - p // 2)])
        err = rng.choice([-eps, eps, 0.0], p) if trial % 3 == 0 else rng.uniform(-eps, eps, p)
        apx = rng.random(p) < 0.9                     # some entries are exact already (multi-candidate pairs)
        e[rng.random(p) < (0.05 if trial % 5 == 0 else 0.0)] = np.nan      # dead pairs
        v = np.where(apx, e + err, e)
        ok = ~np.isnan(v)
        m = int(ok.sum())
        if m == 0:
            continue
        srt = np.sort(v[ok])
        hi = srt[m // 2]
        lo = hi if m % 2 else srt[m // 2 - 1]
        band = 2.0 * eps * (1.0 + 1e-9)
        need = apx & ok & ((np.abs(v - lo) <= band) | (np.abs(v - hi) <= band))
        v2 = np.where(need, e, v)
        assert np.nanmedian(v2) == np.nanmedian(e), (trial, p, kind)
