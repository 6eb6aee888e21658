#!/usr/bin/env bash
# DRAM traffic and per-launch times of every kernel of one bench.py step (run on the GPU box, ONE GPU):
#   scripts/ncu_traffic.sh [workload=c3] [hours=2] [full_hours=24] [extra bench flags...]
# writes gpurun_out/traffic_<key>.csv (ncu launch list), profiles/r2_launches_<key>.csv (per-group summary) and merges
# the per-group DRAM bytes, scaled to the full record length, into profiles/ncu_traffic.json (read by bench.py).
# Recipe: /opt/skills/guides/B200_PROFILING.md (launch-list pass, --clock-control none).  Numbers printed by bench.py
# under ncu are never bench values.
set -euo pipefail
cd "$(dirname "$0")/.."
WL=${1:-c3}; H=${2:-2}; FULL=${3:-24}; shift $(( $# > 3 ? 3 : $# )) || true
KEY="${WL}_${FULL}h_auto"
mkdir -p gpurun_out profiles
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv \
    --log-file "gpurun_out/traffic_${KEY}.csv" \
    python bench.py --workload "$WL" --hours "$H" --steps 1 --warmup 0 --no-cpu --no-e2e --no-profile "$@" \
    > "gpurun_out/traffic_${KEY}.log" 2>&1
python scripts/ncu_traffic.py "gpurun_out/traffic_${KEY}.csv" --key "$KEY" \
    --scale "$(python -c "print($FULL/$H)")" --summary "profiles/r2_launches_${KEY}.csv"
