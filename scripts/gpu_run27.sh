mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python scripts/parity_count.py 3 c3 > gpurun_out/parity_count_c3_3h.json 2> gpurun_out/parity_count.err; python - <<PY
import json
d=json.load(open('gpurun_out/parity_count_c3_3h.json')); d.pop('per_band'); print(d)
PY
timeout 600 python bench.py --steps 6 > gpurun_out/bench_c3_lazy.json 2> gpurun_out/bench_c3_lazy.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_c3_lazy.json') if l.startswith('{')][-1])
print(round(d['value']), d['e2e']['value'], round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()})
print(d['parity'])
print(d['cpu_baseline'])
PY
tail -c 300 gpurun_out/bench_c3_lazy.err
