#!/usr/bin/env bash
# One `ncu --set full` capture of a kernel (run on the GPU box, ONE GPU), summary into profiles/:
#   scripts/ncu_kernel.sh <kernel-name-regex> <tag> [launch-skip=20] [bench flags...]
# e.g. scripts/ncu_kernel.sh xf_pair_kernel r2_xf_pair 40 --hours 1
# Recipe: /opt/skills/guides/B200_PROFILING.md.
set -euo pipefail
cd "$(dirname "$0")/.."
K=$1; TAG=$2; SKIP=${3:-20}; shift $(( $# > 3 ? 3 : $# )) || true
mkdir -p gpurun_out profiles
ncu --set full --clock-control none --import-source on -k "regex:$K" -s "$SKIP" -c 1 -f -o "gpurun_out/$TAG" \
    python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --no-profile "$@" > "gpurun_out/$TAG.log" 2>&1
ncu -i "gpurun_out/$TAG.ncu-rep" --page raw --csv > "gpurun_out/$TAG.raw.csv"
python scripts/ncu_summary.py "gpurun_out/$TAG.raw.csv" > "profiles/$TAG.ncu_summary.txt"
