mkdir -p gpurun_out
for pct in 12 20 30 40; do
timeout 300 python bench.py --opt 9=$pct --no-cpu --no-e2e --no-profile --steps 6 > gpurun_out/bench_c3_ov$pct.json 2> gpurun_out/bench_c3_ov$pct.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_c3_ov$pct.json') if l.startswith('{')][-1])
print('ov$pct', round(d['value']), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()}, round(d['roofline_filter']['frac'],3))
PY
done
