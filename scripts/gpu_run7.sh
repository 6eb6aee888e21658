mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --opt 7=4 --no-cpu --no-e2e > gpurun_out/bench_c3_p1w4.json 2> gpurun_out/bench_c3_p1w4.err
timeout 300 python bench.py --opt 7=8 --no-cpu --no-e2e > gpurun_out/bench_c3_p1w8.json 2> gpurun_out/bench_c3_p1w8.err
timeout 300 python bench.py --opt 7=8 --opt 2=2 --no-cpu --no-e2e > gpurun_out/bench_c3_p1w8_split2.json 2> gpurun_out/bench_c3_p1w8_split2.err
timeout 600 python bench.py --steps 6 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
tail -2 gpurun_out/smoke.log; tail -3 gpurun_out/pytest_gpu.log; for f in bench_c3_p1w4 bench_c3_p1w8 bench_c3_p1w8_split2 bench_c3; do python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
print('$f', round(d['value']), d['stage_ms'], {k:round(v['ms'],2) for k,v in d['kernel_groups'].items() if k.startswith('filter')}, d['e2e'] and (round(d['e2e']['value']), d['e2e']['records_in_flight']))
PY
done; tail -c 300 gpurun_out/*.err
