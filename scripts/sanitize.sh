#!/usr/bin/env bash
# compute-sanitizer memcheck + racecheck on BASELINE config 1 (4 elements, 1 h at 40 sps, 8 bands, OLS) and on a short
# LTS job, through the public Python surface.  Run on the GPU box:  scripts/sanitize.sh
# Logs: profiles/r2_sanitize_{memcheck,racecheck}.log  (exit code != 0 if either tool reports an error).
set -uo pipefail
cd "$(dirname "$0")/.."
mkdir -p profiles
rc=0
for tool in memcheck racecheck; do
    hours=1.0; [ "$tool" = racecheck ] && hours=0.1       # racecheck slows shared-memory kernels by ~100x
    compute-sanitizer --tool "$tool" --error-exitcode 3 --print-limit 20 \
        python scripts/sanitize_job.py "$hours" > "profiles/r2_sanitize_${tool}.log" 2>&1 || rc=$?
    tail -n 4 "profiles/r2_sanitize_${tool}.log"
done
exit $rc
