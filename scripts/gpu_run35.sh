mkdir -p gpurun_out
timeout 200 python bench.py --workload c2 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
timeout 100 python bench.py --workload c1 --no-cpu > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err
python - <<PY
import json
for f in ('bench_c2','bench_c1'):
    d=json.loads([l for l in open('gpurun_out/%s.json'%f) if l.startswith('{')][-1])
    print(f, round(d['value']), round(d['e2e']['value']), round(d['ms_per_step'],3), (d.get('parity') or {}).get('timed_path_vs_cmax_path'))
PY
