#!/usr/bin/env bash
# DRAM traffic and per-launch times of every kernel of one bench.py step (run on the GPU box, ONE GPU):
#   scripts/ncu_traffic.sh [workload=c3] [hours=2] [full_hours=24] [extra bench flags...]
# Two ncu passes (recipe: /opt/skills/guides/B200_PROFILING.md, launch-list pass, --clock-control none):
#   1. every kernel of a step on an `hours`-long record; DRAM bytes of the cross-correlation and solve groups are linear
#      in the record length and are scaled by full_hours/hours (a 24 h step under ncu takes too long);
#   2. the filter kernels alone (-k regex) on the FULL-length record - the share of a chunk their state passes re-read
#      depends on the chunk length, so they do not scale.
# Writes gpurun_out/traffic_<key>{,_filter}.csv (ncu launch lists), gpurun_out/launches_<key>.csv (per-group summary) and
# gpurun_out/ncu_traffic.json; copy the latter two to profiles/ (bench.py reads profiles/ncu_traffic.json).
# Numbers printed by bench.py under ncu are never bench values.
set -euo pipefail
cd "$(dirname "$0")/.."
WL=${1:-c3}; H=${2:-2}; FULL=${3:-24}; shift $(( $# > 3 ? 3 : $# )) || true
KEY="${WL}_${FULL}h_auto"
mkdir -p gpurun_out profiles
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
ncu --metrics $M --clock-control none -c 4000 --csv --log-file "gpurun_out/traffic_${KEY}.csv" \
    python bench.py --workload "$WL" --hours "$H" --steps 1 --warmup 0 --no-cpu --no-e2e --no-profile "$@" \
    > "gpurun_out/traffic_${KEY}.log" 2>&1
python scripts/ncu_traffic.py "gpurun_out/traffic_${KEY}.csv" --key "$KEY" \
    --scale "$(python -c "print($FULL/$H)")" --summary "gpurun_out/launches_${KEY}.csv"
ncu --metrics $M --clock-control none -k "regex:biquad|carry_kernel" -c 64 --csv --log-file "gpurun_out/traffic_${KEY}_filter.csv" \
    python bench.py --workload "$WL" --hours "$FULL" --steps 1 --warmup 0 --no-cpu --no-e2e --no-profile "$@" \
    > "gpurun_out/traffic_${KEY}_filter.log" 2>&1
python scripts/ncu_traffic.py "gpurun_out/traffic_${KEY}_filter.csv" --key "$KEY" --scale 1 --only filter_
cp profiles/ncu_traffic.json gpurun_out/ncu_traffic.json
