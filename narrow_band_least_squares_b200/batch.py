"""Batches of independent arrays (BASELINE config 5: 64 arrays x 24 h, OLS + LTS sweep, sharded by array then segment).

The reference processes several arrays by calling its driver once per array (example_parallel.py:134-138 in a loop;
narrow_band_least_squares.py:223-323); nothing is shared between arrays.  Here:

  * ``deal_arrays``  arrays are dealt to the ranks of a torchrun job FIRST (array a -> rank a mod world); only when
                     there are fewer arrays than ranks does a group of ranks share one array by TIME SEGMENT with filter
                     halos (sharding.Shard) — SURVEY.md §8(e).  No data-path collective either way.
  * ``narrow_band_least_squares_batch``  on one GPU the arrays of a rank go through ``in_flight`` libnbls contexts
                     (own stream and workspace each, one host thread each), so the H2D copy and host-side assembly of
                     one array overlap the kernels of another.
  * OLS + LTS sweep  ``alphas=(1.0, 0.75)``: the band-pass and the cross-correlation (> 90 % of a run) are done once per
                     array; every further alpha re-solves on the lags the context kept (nbls_resolve_host).

Each array's result is bit-identical to a stand-alone ``narrow_band_least_squares`` call with the same arguments
(tests/test_gpu_wide.py::test_batch_equals_standalone_runs).
"""
from __future__ import annotations

import queue
import threading

import numpy as np

from . import _capi
from .engine import Engine, Results
from .narrow_band_least_squares import assemble_outputs, prepare_job


def deal_arrays(narrays, world):
    """Work items per rank: ``items[rank] = [(array index, segment rank, segment world), ...]``.

    Arrays first: with narrays >= world, array a goes whole to rank a mod world.  With fewer arrays than ranks, the
    ranks are split into narrays groups and each group shares its array by time segment.
    """
    items = [[] for _ in range(world)]
    if narrays >= world:
        for a in range(narrays):
            items[a % world].append((a, 0, 1))
        return items
    base, extra = divmod(world, narrays)
    r = 0
    for a in range(narrays):
        g = base + (1 if a < extra else 0)
        for s in range(g):
            items[r].append((a, s, g))
            r += 1
    return items


def narrow_band_least_squares_batch(records, WINLEN_list, WINOVER, ALPHA, NBANDS, w, h, freqlist, FREQ_BAND_TYPE,
                                    freq_resp_list, FILTER_TYPE, FILTER_ORDER, FILTER_RIPPLE, *, alphas=None,
                                    device=None, engines=None, in_flight=2, max_lag=-1, method=_capi.XCORR_AUTO,
                                    rank=0, world=1):
    """``narrow_band_least_squares`` over a list of independent arrays.

    ``records``: sequence of ``(st, lat_list, lon_list)`` (or ``(st, None, None, rij)`` with (2, C) km coordinates);
    the remaining positional arguments are the reference driver's (narrow_band_least_squares.py:8), shared by all arrays.
    ``alphas``: sweep — a sequence of ALPHA values solved on ONE filter + cross-correlation pass per array; default
    ``(ALPHA,)``.  Under torchrun pass ``rank`` / ``world``: this rank processes the arrays ``deal_arrays`` gives it
    (whole arrays; time-segment groups are for callers that drive ``sharding.Shard`` themselves, e.g. bench.py).

    Returns a list with one entry per array of THIS rank, ``(array index, {alpha: 9-tuple})`` — each 9-tuple is what
    the reference driver returns for that array and ALPHA.
    """
    alphas = tuple(alphas) if alphas is not None else (ALPHA,)
    mine = [a for a, s, g in deal_arrays(len(records), world)[rank] if s == 0]
    if engines is None:
        import os
        dev = int(os.environ.get("LOCAL_RANK", "0")) if device is None else int(device)
        engines = [Engine(dev) for _ in range(max(1, min(in_flight, len(mine))))]
        own = True
    else:
        own = False
    todo = queue.Queue()
    for a in mine:
        todo.put(a)
    out, errors, lock = {}, [], threading.Lock()

    def worker(eng):
        try:
            while True:
                try:
                    a = todo.get_nowait()
                except queue.Empty:
                    return
                rec = records[a]
                st, lat, lon = rec[0], rec[1], rec[2]
                rij = rec[3] if len(rec) > 3 else None
                data, tvec, co, plan = prepare_job(WINLEN_list, WINOVER, alphas[0], st, lat, lon, NBANDS, freqlist,
                                                   FREQ_BAND_TYPE, FILTER_TYPE, FILTER_ORDER, FILTER_RIPPLE, rij=rij,
                                                   max_lag=max_lag, method=method)
                per_alpha = {}
                for k, alpha in enumerate(alphas):
                    p = plan if k == 0 else plan.with_alpha(alpha)
                    res = Results(p, pinned=False, want_cmax=False)
                    if k == 0:
                        eng.run(p, data, res)
                    else:
                        eng.resolve(p, res)
                    per_alpha[alpha] = assemble_outputs(p, res, tvec, co, WINLEN_list, WINOVER, alpha, NBANDS, w, h,
                                                        freqlist, FREQ_BAND_TYPE, freq_resp_list, FILTER_TYPE,
                                                        FILTER_ORDER, FILTER_RIPPLE)
                with lock:
                    out[a] = per_alpha
        except BaseException as exc:      # noqa: BLE001 - re-raised below
            errors.append(exc)

    threads = [threading.Thread(target=worker, args=(e,)) for e in engines]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if own:
        for e in engines:
            e.close()
    if errors:
        raise errors[0]
    return [(a, out[a]) for a in mine]
