mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --no-cpu --steps 3 > gpurun_out/bench_last.json 2> gpurun_out/bench_last.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_last.json') if l.startswith('{')][-1])
print(round(d['value']), round(d['e2e']['value']), round(d['ms_per_step'],2), d['clocks']['reasons'])
PY
