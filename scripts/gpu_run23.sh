mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 400 python bench.py --no-cpu --no-e2e --steps 6 > gpurun_out/bench_c3_ld8.json 2> gpurun_out/bench_c3_ld8.err
tail -4 gpurun_out/pytest_gpu.log; python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_c3_ld8.json') if l.startswith('{')][-1])
print(round(d['value']), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()})
print({k:round(v['ms'],2) for k,v in d['kernel_groups'].items()} if isinstance(d.get('kernel_groups'),dict) else d.get('kernel_groups'))
PY
tail -c 300 gpurun_out/bench_c3_ld8.err
