mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --opt 9=0 --no-cpu --no-e2e --steps 6 > gpurun_out/bench_c3_ov0.json 2> gpurun_out/bench_c3_ov0.err
timeout 300 python bench.py --opt 9=1 --no-cpu --no-e2e --steps 6 > gpurun_out/bench_c3_ov1.json 2> gpurun_out/bench_c3_ov1.err
timeout 300 python bench.py --workload c4 --hours 72 --opt 9=0 --no-cpu --no-e2e --steps 2 --warmup 1 > gpurun_out/bench_c4_ov0.json 2> gpurun_out/bench_c4_ov0.err
timeout 300 python bench.py --workload c4 --hours 72 --opt 9=1 --no-cpu --no-e2e --steps 2 --warmup 1 > gpurun_out/bench_c4_ov1.json 2> gpurun_out/bench_c4_ov1.err
tail -4 gpurun_out/pytest_gpu.log; for f in bench_c3_ov0 bench_c3_ov1 bench_c4_ov0 bench_c4_ov1; do python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
print('$f', round(d['value']), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()}, round(d['roofline_filter']['frac'],3))
PY
done; tail -c 300 gpurun_out/*.err
