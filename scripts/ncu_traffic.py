#!/usr/bin/env python
"""Turn an ncu CSV launch list (metrics dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum) of one
bench.py step into DRAM bytes per kernel group per step and merge them into profiles/ncu_traffic.json, the file
bench.py reads `traffic` from.

  python scripts/ncu_traffic.py gpurun_out/traffic.csv --key c3_24h_auto --scale 12 [--summary profiles/r2_x.csv]

--scale: the capture ran on a record 1/scale of the workload's length (ncu replays every launch, a 24 h step under ncu
takes too long); traffic of every group is linear in the record length, so the sums are multiplied by it.
"""
import argparse
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def group_of(name, state):
    if "fma_peak_kernel" in name:
        return None
    if "biquad_state_shared_kernel" in name:
        return "filter_fwd_phase1"
    if "carry_kernel" in name:
        return "filter_carry"
    if "biquad_chunk_kernel" in name:
        rev = re.search(r"biquad_chunk_kernel<[^,>]+,\s*(?:\(bool\))?\s*(true|1)", name) is not None
        k = "rev" if rev else "fwd"
        state[k] = state.get(k, 0) + 1
        if rev:      # backward pass: phase 1 then phase 3, alternating
            return "filter_bwd_phase1" if state[k] % 2 == 1 else "filter_bwd_phase3"
        return "filter_fwd_phase3" if state.get("shared_p1", True) else ("filter_fwd_phase1" if state[k] % 2 == 1 else "filter_fwd_phase3")
    for pat, g in (("xf_spectra_kernel", "xf_spectra"), ("xf_pair_kernel", "xf_pair"), ("xf_verify_items_kernel", "xf_verify_items"), ("xf_median_select_kernel", "xf_verify_items"),
                   ("xf_verify_kernel", "xf_verify"), ("xcorr_direct_kernel", "xcorr_direct"), ("stats_ols_kernel", "stats_ols"),
                   ("lts_kernel", "lts"), ("xf_audit_kernel", None)):
        if pat in name:
            return g
    return "other"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--key", required=True)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--summary", default=None, help="also write a per-group summary CSV here")
    ap.add_argument("--only", default=None, help="merge only the groups whose name starts with this prefix into the "
                    "existing entry (e.g. 'filter_': the filter's state passes are not linear in the record length, "
                    "so they are captured at full length in a pass of their own)")
    a = ap.parse_args()
    rows = {}
    with open(a.csv, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        rid = int(r["ID"])
        d = rows.setdefault(rid, {"name": r["Kernel Name"]})
        try:
            val = float(r["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        unit = r.get("Metric Unit", "")
        m = r["Metric Name"]
        if m.startswith("dram__bytes"):
            mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1.0)
            d[m] = val * mult
        elif m.startswith("gpu__time_duration"):
            mult = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}.get(unit, 1e-6)
            d["ms"] = val * mult
    state, groups = {}, {}
    state["shared_p1"] = any("biquad_state_shared_kernel" in d["name"] for d in rows.values())
    for rid in sorted(rows):
        d = rows[rid]
        g = group_of(d["name"], state)
        if g is None:
            continue
        e = groups.setdefault(g, {"launches": 0, "bytes": 0.0, "ms": 0.0})
        e["launches"] += 1
        e["bytes"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        e["ms"] += d.get("ms", 0.0)
    out_path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(out_path) as f:
            allk = json.load(f)
    except Exception:
        allk = {}
    if a.only:
        groups = {g: e for g, e in groups.items() if g.startswith(a.only)}
        entry, meta = allk.get(a.key, {}), allk.get(a.key + "__meta", {})
        entry.update({g: e["bytes"] * a.scale for g, e in groups.items()})
        meta.setdefault("passes", []).append({"only": a.only, "scale": a.scale, "source_csv": os.path.basename(a.csv),
                                              "ncu_ms_unscaled": {g: e["ms"] for g, e in groups.items()}})
        allk[a.key], allk[a.key + "__meta"] = entry, meta
    else:
        allk[a.key] = {g: e["bytes"] * a.scale for g, e in groups.items()}
        allk[a.key + "__meta"] = {"scale": a.scale, "source_csv": os.path.basename(a.csv),
                                  "launches": {g: e["launches"] for g, e in groups.items()},
                                  "ncu_ms_unscaled": {g: e["ms"] for g, e in groups.items()}}
    with open(out_path, "w") as f:
        json.dump(allk, f, indent=1, sort_keys=True)
    if a.summary:
        tot = sum(e["ms"] for e in groups.values()) or 1.0
        with open(a.summary, "w") as f:
            f.write("group,launches,ncu_ms,share,dram_GB,dram_GB_scaled\n")
            for g, e in sorted(groups.items(), key=lambda kv: -kv[1]["ms"]):
                f.write(f"{g},{e['launches']},{e['ms']:.3f},{e['ms'] / tot:.4f},{e['bytes'] / 1e9:.3f},{e['bytes'] * a.scale / 1e9:.3f}\n")
    print(json.dumps(allk[a.key]), file=sys.stderr)


if __name__ == "__main__":
    main()
