mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for sub in 1 2 4 8; do
timeout 300 python bench.py --opt 8=$sub --no-cpu --no-e2e > gpurun_out/bench_c3_sub$sub.json 2> gpurun_out/bench_c3_sub$sub.err
done
tail -4 gpurun_out/pytest_gpu.log; for f in bench_c3_sub1 bench_c3_sub2 bench_c3_sub4 bench_c3_sub8; do python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
print('$f', round(d['value']), round(d['stage_ms']['filter'],2), {k:round(v['ms'],2) for k,v in d['kernel_groups'].items() if k.startswith('filter')}, round(d['roofline_filter']['frac'],3))
PY
done; tail -c 300 gpurun_out/*.err
