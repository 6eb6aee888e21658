mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 6 > gpurun_out/bench_c3_final4.json 2> gpurun_out/bench_c3_final4.err; tail -c 300 gpurun_out/bench_c3_final4.err
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
python - <<PY
import json
for f in ('bench_c3_final4','bench_default'):
    d=json.loads([l for l in open('gpurun_out/%s.json'%f) if l.startswith('{')][-1])
    print(f, round(d['value']), round(d['e2e']['value']), d['steps'], round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()}, d['gpu_launches'], d['clocks']['reasons'], round(d['roofline']['frac'],3), d['xcorr_stats']['verified'], d['xcorr_stats']['max_screen_err'])
    print({k:round(v['ms'],2) for k,v in d['kernel_groups'].items()})
    print(d['parity']['timed_path_vs_cmax_path'], d['parity']['vs_oracle']['lag_mismatches'], d['cpu_baseline']['value'])
PY
