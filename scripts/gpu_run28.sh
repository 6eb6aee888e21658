mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/$name.json') if l.startswith('{')][-1])
    print('$name', round(d['value']), 'e2e', d.get('e2e',{}).get('value') if d.get('e2e') else None, round(d['ms_per_step'],2), d.get('parity',{}).get('vs_oracle',{}).get('lag_mismatches') if d.get('parity') else None)
except Exception as e:
    print('$name FAILED', e); print(open('gpurun_out/$name.err').read()[-600:])
PY
}
run bench_c3 --steps 6
run bench_c2 --workload c2
run bench_c1 --workload c1
run bench_c4_720h --workload c4 --steps 1 --warmup 1 --e2e-in-flight 1 --no-cpu
run bench_c5_64 --workload c5 --steps 1 --warmup 1 --no-cpu
run bench_c3_maxlag200 --max-lag 200 --no-cpu
run bench_c3_single_stage --single-stage --no-cpu --no-e2e
run bench_c3_noise_only --noise-only --no-cpu --no-e2e
run bench_c3_lazy_off --opt 10=0 --no-cpu --no-e2e
