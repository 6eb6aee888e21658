mkdir -p gpurun_out
timeout 120 narrow_band_least_squares_b200/_lib/fp32x2_microbench > gpurun_out/fp32x2_microbench.jsonl; cat gpurun_out/fp32x2_microbench.jsonl
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; tail -c 400 gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_reference_arm.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_default.json') if l.startswith('{')][-1])
print(round(d['value']), d['e2e']['value'], d['steps'], d['warmup'], d['gpu_launches'], d['clocks'], d['roofline']['frac'], d['roofline']['xcorr_stage'])
print(d['parity']['timed_path_vs_cmax_path'], d['parity']['vs_oracle']['lag_mismatches'])
r=json.loads([l for l in open('gpurun_out/bench_reference_arm.json') if l.startswith('{')][-1]); print({k:r[k] for k in ('impl','value','unit','ms_per_step')}, r['cpu_baseline']['sample'][:80])
PY
