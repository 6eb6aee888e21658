mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 6 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
timeout 300 python bench.py --workload c2 --no-cpu > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
tail -4 gpurun_out/pytest_gpu.log; for f in bench_c3 bench_c2; do python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
print('$f', round(d['value']), {k:round(v,2) for k,v in d['stage_ms'].items()}, {k:round(v['ms'],2) for k,v in d['kernel_groups'].items()}, d['e2e'] and (round(d['e2e']['value']), d['e2e']['records_in_flight']), d.get('parity',{}).get('vs_oracle'))
PY
done; tail -c 300 gpurun_out/*.err
