mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --no-cpu --no-e2e --steps 6 > gpurun_out/bench_c3_rmw.json 2> gpurun_out/bench_c3_rmw.err
tail -4 gpurun_out/pytest_gpu.log; python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_c3_rmw.json') if l.startswith('{')][-1])
print('rmw', round(d['value']), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()}, round(d['roofline_filter']['frac'],3), {k:round(v['ms'],2) for k,v in d['kernel_groups'].items() if k.startswith('filter')})
PY
tail -c 300 gpurun_out/bench_c3_rmw.err
