#!/usr/bin/env python
"""bench.py — NBLS band-window estimates per second on B200 (BASELINE.json metric).

  python bench.py [--gpus N --steps K --warmup W] [--impl reference] [--workload c3] [--hours H]

One "step" = one pass of the whole hot path (per-band zero-phase band-pass -> windowed pairwise
cross-correlation -> slowness solve) over one synthetic record of the workload (default c3: 25
elements, 24 h at 100 sps, 20 log bands 0.5-5 Hz, OLS, adaptive 60->30 s windows, 50 % overlap).
N > 1 (torchrun): the record is N x 24 h long and is sharded by time segment with filter halos,
one rank per GPU (weak scaling); the per-window result arrays are gathered with NCCL.

`value`   = estimates / device time with the record resident in HBM (CUDA events on libnbls' stream).
`e2e`     = the same through the host-buffer C-ABI call nbls_run_host (pinned H2D + path + D2H), by default with two
            records in flight on two contexts (`--e2e-in-flight 1` = strictly one after the other).
`roofline`= the dominant kernel (cross-correlation, FP64-FMA bound) against the DFMA peak measured
            in this run; `roofline_filter` = the HBM-bound filter kernels against MEASURED_PEAKS.json.
`cpu_baseline` / --impl reference = the reference's algorithm (oracle port: the same numpy/scipy calls,
            joblib over bands like example_parallel.py) on the box's host cores, bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FALLBACK_HBM_GBS = 6650.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--hours", type=float, default=None, help="record length per GPU (default: the workload's)")
    ap.add_argument("--method", default="auto", choices=["auto", "direct", "fft", "fft64", "fft32"])
    ap.add_argument("--max-lag", type=int, default=-1, help="restrict |lag| (default: all 2N-1 lags = reference)")
    ap.add_argument("--cpu-sample-min", type=float, default=3.0, help="minutes of record for the CPU legs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-in-flight", type=int, default=2,
                    help="records in flight in the e2e leg (own context each; 1 = strictly one after the other)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi sampling during the timed region (B200_PROFILING.md clocks line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.t_mark = index, [], None, None

    def mark(self):
        """Start of the timed region: only samples read after this count (nvidia-smi is started before the warm-up
        steps so that its start-up does not stall the first timed step)."""
        self.t_mark = time.perf_counter()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        rows = [r for t, r in self.rows if self.t_mark is None or t >= self.t_mark]
        if not rows:                      # timed region shorter than one sampling period: take the last sample
            rows = [r for _, r in self.rows[-1:]]
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
            except Exception:
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7),
                              ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": FALLBACK_HBM_GBS}, "fallback"


def workload(args):
    """Host-side description of the job: config dict, per-GPU npts, record generator."""
    from narrow_band_least_squares_b200.configs import config_job
    j = config_job(args.workload)
    hours = j["hours"] if args.hours is None else args.hours
    j["hours_per_gpu"] = hours
    j["npts_per_gpu"] = int(round(hours * 3600.0 * j["fs"]))
    return j


def make_record(j, npts, seed_offset=0):
    from narrow_band_least_squares_b200.synth import CONFIGS, plane_wave_record
    c = CONFIGS[j["name"]]
    return plane_wave_record(c["seed"] + seed_offset, c["nchans"], npts, c["fs"], c["aperture_km"],
                             flip_channel=c.get("flip_channel"), shift_channel=c.get("shift_channel"))


def run_b200(args):
    import ctypes as C

    from narrow_band_least_squares_b200 import _capi
    from narrow_band_least_squares_b200.engine import DeviceArray, Engine, Plan, Results
    from narrow_band_least_squares_b200.engine import window_grid
    from narrow_band_least_squares_b200.helpers import band_edges, design_sos
    from narrow_band_least_squares_b200.sharding import Shard, filter_halo, gather_results

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    j = workload(args)
    method = {"auto": _capi.XCORR_AUTO, "direct": _capi.XCORR_DIRECT, "fft": _capi.XCORR_FFT, "fft64": _capi.XCORR_FFT,
              "fft32": _capi.XCORR_FFT32}[args.method]
    fs, nchan, nb = j["fs"], j["nchans"], j["nbands"]
    g_npts = j["npts_per_gpu"] * world                       # weak scaling: N x the record length
    sos = np.stack([design_sos(j["filter_type"], *band_edges(j["freqlist"], b, "log"), fs, j["filter_order"],
                               j["filter_ripple"]) for b in range(nb)])
    grids = [window_grid(g_npts, fs, j["winlen_list"][b], j["winover"]) for b in range(nb)]
    win_len = [g[0] for g in grids]; win_hop = [g[1] for g in grids]; win_cnt = [g[2] for g in grids]
    halo = filter_halo(sos) if world > 1 else 0
    shard = Shard(g_npts, win_len, win_hop, win_cnt, rank, world, halo)
    rec = make_record(j, shard.npts, seed_offset=rank)      # each rank synthesises its own segment
    plan = shard.plan(nchan, fs, sos, j["filter_type"] == "butter", win_len, win_hop, rec.rij, alpha=j["alpha"],
                      max_lag=args.max_lag, method=method)
    e_local, e_total = plan.nwin, int(sum(win_cnt))
    eng = Engine(local_rank)
    ld = (shard.npts + 15) // 16 * 16

    # pinned host input + results (the e2e leg copies from / to these every step)
    x_host = _capi.pinned_empty((nchan, shard.npts))
    x_host[:] = rec.data
    res = Results(plan, pinned=True, want_cmax=False)
    # device-resident copy for the `value` leg
    xpad = np.zeros((nchan, ld))
    xpad[:, :shard.npts] = rec.data
    x_dev = DeviceArray(eng, xpad.nbytes).upload(xpad)
    del xpad
    if world > 1:
        import torch
        t_out = {k: torch.empty(max(e_local, 1), dtype=torch.float64, device="cuda") for k in
                 ("mdccm", "mad", "vel", "baz", "sigma_tau")}
        t_out["slowness"] = torch.empty((max(e_local, 1), 2), dtype=torch.float64, device="cuda")
        t_out["lag"] = torch.empty((max(e_local, 1), plan.npairs), dtype=torch.int32, device="cuda")
        if plan.alpha != 1.0:
            t_out["weights"] = torch.empty((max(e_local, 1), plan.npairs), dtype=torch.uint8, device="cuda")
        o_dev = _capi.Outputs()
        for k, t in t_out.items():
            setattr(o_dev, k, t.data_ptr())
        keep = None
    else:
        o_dev, keep = eng.device_outputs(plan)
        o_dev.cmax = None

    def barrier():
        eng.synchronize()
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()

    def step_resident():
        eng.run_resident(plan, x_dev, ld, o_dev)
        eng.synchronize()
        t = eng.timings()
        ms_gather = 0.0
        if dist is not None:
            import torch
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            gather_results({k: v[:e_local] for k, v in t_out.items() if k != "lag"}, plan.win_count)
            e1.record()
            torch.cuda.synchronize()
            ms_gather = e0.elapsed_time(e1)
        return t, ms_gather

    fp64_peak = eng.fp64_peak_tflops(20000)
    fp32_peak = eng.fp32_peak_tflops(20000)
    sampler = ClockSampler(local_rank).start()
    for _ in range(args.warmup):
        step_resident()
    barrier()
    launches0 = eng.launch_count()
    eng.xcorr_stats(reset=True)
    sampler.mark()
    wall0 = time.perf_counter()
    dev_ms, stage = 0.0, {"filter": 0.0, "xcorr": 0.0, "solve": 0.0}
    for _ in range(args.steps):
        t, mg = step_resident()
        dev_ms += t["total"] + mg
        for k in stage:
            stage[k] += t[k]
    barrier()
    wall_ms = (time.perf_counter() - wall0) * 1e3
    clocks = sampler.stop()
    launches = eng.launch_count() - launches0
    xstats = eng.xcorr_stats()
    xstats["pair_windows"] = int(sum(int(w) for w in plan.win_count) * plan.npairs * args.steps)
    if dist is not None:
        import torch
        tt = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dev_ms, wall_ms = float(tt[0]), float(tt[1])
    ms_per_step = dev_ms / args.steps
    value = e_total / (ms_per_step * 1e-3)

    # ---- e2e: host buffers through nbls_run_host, every step H2D + D2H ----
    # `--e2e-in-flight n` (default 2) keeps n records in flight, each on its own context (own stream and workspace): the
    # H2D copy and the host-side waits of record k+1 overlap the kernels of record k, as a caller working through a
    # queue of records would run it.  Every step still copies its input from pinned host memory and its results back.
    e2e = None
    if not args.no_e2e:
        import threading
        nfl = max(1, min(args.e2e_in_flight, args.steps))
        engs = [eng] + [Engine(local_rank) for _ in range(nfl - 1)]
        ress = [res] + [Results(plan, pinned=True, want_cmax=False) for _ in range(nfl - 1)]
        for e_, r_ in zip(engs, ress):
            e_.run(plan, x_host, r_)                          # one warm-up each (buffers grow once)
        for r_ in ress[1:]:                                   # every context computes the same bits
            assert np.array_equal(r_.lag, res.lag) and np.array_equal(r_.vel, res.vel, equal_nan=True)
        barrier()
        done = [threading.Event() for _ in range(args.steps)]
        gathered = [threading.Event() for _ in range(args.steps)]
        step_ms = [0.0] * args.steps
        errors = []

        def worker(i):
            try:
                for k in range(i, args.steps, nfl):
                    if k - nfl >= 0:
                        gathered[k - nfl].wait()              # the results buffer of this context is free again
                    engs[i].run(plan, x_host, ress[i])
                    step_ms[k] = engs[i].timings()["total"]
                    done[k].set()
            except BaseException as exc:                      # noqa: BLE001 - re-raised in the main thread
                errors.append(exc)
                for ev in done:
                    ev.set()

        w0 = time.perf_counter()
        threads = [threading.Thread(target=worker, args=(i,)) for i in range(nfl)]
        for th in threads:
            th.start()
        for k in range(args.steps):
            done[k].wait()
            if errors:
                break
            if dist is not None:
                r_ = ress[k % nfl]
                gather_results({kk: getattr(r_, kk)[:e_local] for kk in ("mdccm", "vel", "baz", "sigma_tau")},
                               plan.win_count)
            gathered[k].set()
        for ev in gathered:
            ev.set()
        for th in threads:
            th.join()
        if errors:
            raise errors[0]
        barrier()
        e_wall = (time.perf_counter() - w0) * 1e3
        e_ms = sum(step_ms)
        if dist is not None:
            import torch
            tt = torch.tensor([e_wall], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            e_wall = float(tt[0])
        e2e = {"value": e_total / (e_wall / args.steps * 1e-3), "unit": "estimates/s",
               "h2d_bytes_per_step": int(x_host.nbytes) * world, "d2h_bytes_per_step": int(res.nbytes()) * world,
               "ms_per_step_wall": e_wall / args.steps, "ms_record_latency_device": e_ms / args.steps,
               "records_in_flight": nfl}
    return dict(j=j, plan=plan, world=world, rank=rank, value=value, ms_per_step=ms_per_step, wall_ms=wall_ms,
                stage={k: v / args.steps for k, v in stage.items()}, clocks=clocks, launches=launches, e2e=e2e,
                fp64_peak=fp64_peak, fp32_peak=fp32_peak, xstats=xstats, e_total=e_total, e_local=e_local, shard=shard, sos=sos, win_len=win_len,
                win_cnt=win_cnt, g_npts=g_npts, dist=dist, keep=keep)


def cpu_port_rate(j, minutes, n_jobs):
    """The reference's algorithm on the host cores (oracle port; joblib over bands like
    narrow_band_least_squares_parallel, reference :285) on a `minutes`-long slice of the workload."""
    from oracle import nbls_oracle as orc
    npts = int(round(minutes * 60.0 * j["fs"]))
    rec = make_record(j, npts)
    tvec = np.arange(npts) / j["fs"] / 86400.0
    t0 = time.perf_counter()
    out = orc.nbls_oracle(rec.data, j["fs"], tvec, rec.rij, j["freqlist"], j["winlen_list"], j["winover"], j["alpha"],
                          ftype=j["filter_type"], order=j["filter_order"], ripple=j["filter_ripple"], n_jobs=n_jobs)
    dt = time.perf_counter() - t0
    est = int(sum(out["num_compute"]))
    return est / dt, est, dt


def config_block(j, world, args, extra=None):
    c = {"workload": f"{j['name']}: {j['nchans']}-element array, {j['hours_per_gpu']:g} h at {j['fs']:g} sps per GPU, "
                     f"{j['nbands']} log bands 0.5-5 Hz, alpha={j['alpha']:g}, windows {j['winlen_list'][0]:g}->"
                     f"{j['winlen_list'][-1]:g} s, 50% overlap, butter order {j['filter_order']} zero-phase",
         "sharding": "1 GPU" if world == 1 else f"time segments with filter halos over {world} GPUs, record = {world} x {j['hours_per_gpu']:g} h",
         "l2": "inputs larger than L2 (record and filtered bands are GB-sized)", "xcorr_method": args.method,
         "max_lag": args.max_lag,
         "e2e": (f"{max(1, min(args.e2e_in_flight, args.steps))} record(s) in flight, one context each; every step copies its "
                 "input from pinned host memory and its results back inside the timed region")}
    if extra:
        c.update(extra)
    return c


def main():
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    ncores = os.cpu_count() or 1
    if args.impl == "reference":
        if rank != 0:
            return
        j = workload(args)
        for _ in range(max(0, min(args.warmup, 1))):          # numba JIT etc. (LTS workloads)
            cpu_port_rate(j, min(args.cpu_sample_min, 2.0), ncores)
        rates, est, secs = [], 0, 0.0
        for _ in range(args.steps):
            r, est, dt = cpu_port_rate(j, args.cpu_sample_min, ncores)
            rates.append(r); secs += dt
        v = float(np.mean(rates))
        sample = f"{args.cpu_sample_min:g}-minute slice of the workload, all {j['nbands']} bands, {est} estimates per step"
        print(json.dumps({"impl": "reference", "metric": "NBLS band-window estimates/sec", "value": v,
                          "unit": "estimates/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                          "config": config_block(j, 1, args),
                          "cpu_baseline": {"value": v, "unit": "estimates/s", "cores": ncores, "kind": "port",
                                           "sample": sample},
                          "e2e": {"value": v, "unit": "estimates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    r = run_b200(args)
    if r["rank"] != 0:
        if r["dist"] is not None:
            r["dist"].destroy_process_group()
        return
    j, plan = r["j"], r["plan"]
    peaks, peak_src = measured_peaks()
    # algorithmic work of this rank's launches (formulas restated in DESIGN.md §Measurement)
    from narrow_band_least_squares_b200 import _capi
    lib = _capi.load()
    B, C, P = plan.nbands, plan.nchan, plan.npairs
    T = plan.npts
    fp32 = args.method in ("auto", "fft32")
    flop, flop64, n_fft_bands = 0.0, 0.0, 0
    for b in range(B):
        W, N = int(plan.win_count[b]), int(plan.win_len[b])
        L = 0
        if args.method != "direct" and (args.method in ("fft", "fft64", "fft32") or N >= 128):
            L = int(lib.nbls_xcorr_fft_length(N, args.max_lag))
        if L:      # FFT path: ceil(C/2) forward + ceil(P/2) inverse complex FFTs of L points (5 L log2 L each),
            n_fft_bands += 1       # cross-spectra 2 x 6 flop and screening/arg-max 2 compares per point of the inverse
            flop += W * (((C + 1) // 2 + (P + 1) // 2) * 5.0 * L * np.log2(L) + ((P + 1) // 2) * L * 16.0)
            if fp32:   # exact FP64 evaluation of ~1 candidate lag per pair: N multiply-adds
                flop64 += W * P * 2.0 * N
        else:      # time-domain path: M^2 multiply-adds per pair, M = N rounded up to 8
            flop64 += W * P * 2.0 * (-(-N // 8) * 8.0) ** 2
    x_ms, f_ms = r["stage"]["xcorr"], r["stage"]["filter"]
    use32 = fp32 and n_fft_bands > 0
    work = flop if n_fft_bands > 0 else flop64
    peak = r["fp32_peak"] if use32 else r["fp64_peak"]
    tfl = work / (x_ms * 1e-3) / 1e12 if x_ms > 0 else None
    filt_bytes = 8.0 * C * T * (1 + 3 * B)
    fgb = filt_bytes / (f_ms * 1e-3) / 1e9 if f_ms > 0 else None
    if n_fft_bands == 0:
        kern = "xcorr_direct_kernel"
    elif use32:
        kern = "xf_pair_kernel + xf_spectra_kernel (FP32 screening FFTs) + xf_verify_kernel (exact FP64 candidates)"
    else:
        kern = "xf_pair_kernel + xf_spectra_kernel (FP64 FFT cross-correlation)"
    if n_fft_bands not in (0, B):
        kern += " + xcorr_direct_kernel on the short-window bands"
    line = {"metric": "NBLS band-window estimates/sec", "value": r["value"], "unit": "estimates/s",
            "n_gpus": r["world"], "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_block(j, r["world"], args, {"estimates_per_step": r["e_total"]}),
            "clocks": r["clocks"], "gpu_launches": int(r["launches"]), "e2e": r["e2e"],
            "stage_ms": r["stage"], "wall_ms_per_step": r["wall_ms"] / args.steps,
            "roofline": {"kernel": kern, "bound": "fp32" if use32 else "fp64", "achieved": tfl, "peak": peak,
                         "unit": "TFLOP/s", "frac": (tfl / peak) if tfl else None,
                         # DRAM bytes of all cross-correlation kernels per step, ncu (profiles/r1_xcorr_dram_traffic_2h.csv
                         # x 12): 0.4 TB/s over the stage, i.e. not the bound
                         "traffic": 184.7e9 if (args.workload == "c3" and args.hours is None and r["world"] == 1
                                                and args.method == "auto") else None,
                         "peak_source": ("FP32" if use32 else "FP64") + " FMA microbenchmark run in this process, 2 flop per "
                                        "FMA (MEASURED_PEAKS.json has no such entry; the path is neither HBM- nor tensor-"
                                        "bound).  FFT butterflies are mostly unfused adds, so 0.5 is the ceiling of this "
                                        "fraction; ncu (profiles/) gives the issue-slot utilisation of the pair kernel",
                         "algorithmic_flop_per_step": work, "fp64_verify_flop_per_step": flop64 if use32 else None,
                         "note": ("algorithmic = the single-stage search over all 2N-1 lags; with --method auto the "
                                  "two-stage search executes less (xcorr_stats.two_stage_*), so 'achieved' is "
                                  "algorithmic work per second, not executed flop/s") if args.method == "auto" else None,
                         "fp64_peak": r["fp64_peak"], "fp32_peak": r["fp32_peak"]},
            "roofline_filter": {"kernel": "biquad_state_shared_kernel + biquad_chunk_kernel x3 + carry_kernel x2", "bound": "hbm", "achieved": fgb,
                                "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": (fgb / peaks["hbm_gbs"]) if fgb else None,
                                "peak_source": peak_src, "algorithmic_bytes": filt_bytes,
                                # dram__bytes_read + write of the six filter kernels, ncu capture of this workload at its
                                # default length (profiles/r1_filter_24h_kernels.csv); not measured live
                                "traffic": 122.8e9 if (args.workload == "c3" and args.hours is None and r["world"] == 1)
                                else None}}
    try:
        st = r["xstats"]
        line["xcorr_stats"] = st
    except KeyError:
        pass
    if not args.no_cpu:
        v, est, dt = cpu_port_rate(j, args.cpu_sample_min, ncores)
        line["cpu_baseline"] = {"value": v, "unit": "estimates/s", "cores": ncores, "kind": "port",
                                "sample": f"{args.cpu_sample_min:g}-minute slice, all {j['nbands']} bands, {est} estimates in {dt:.1f} s"}
    print(json.dumps(line))
    if r["dist"] is not None:
        r["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
