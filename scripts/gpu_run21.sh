mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 400 python bench.py --workload c4 --hours 72 --steps 2 --warmup 1 --e2e-in-flight 1 --no-cpu > gpurun_out/bench_c4_72h.json 2> gpurun_out/bench_c4_72h.err
timeout 300 python bench.py --workload c2 --no-cpu > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
tail -6 gpurun_out/pytest_gpu.log; for f in bench_c4_72h bench_c2; do python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
print('$f', round(d['value']), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()})
PY
done; tail -c 300 gpurun_out/bench_c4_72h.err
