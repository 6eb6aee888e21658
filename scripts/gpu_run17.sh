mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 6 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
timeout 300 python bench.py --workload c2 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
timeout 300 python bench.py --workload c1 > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err
timeout 300 python bench.py --max-lag 200 --no-cpu > gpurun_out/bench_c3_maxlag200.json 2> gpurun_out/bench_c3_maxlag200.err
timeout 600 bash scripts/ncu_traffic.sh c3 2 24 > gpurun_out/ncu_traffic.out 2>&1
tail -2 gpurun_out/smoke.log; tail -3 gpurun_out/pytest_gpu.log; for f in bench_c3 bench_c2 bench_c1 bench_c3_maxlag200; do python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
print('$f', round(d['value']), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()}, round(d['roofline_filter']['frac'],3), d['e2e'] and (round(d['e2e']['value']), d['e2e']['records_in_flight']))
PY
done; tail -c 300 gpurun_out/*.err
