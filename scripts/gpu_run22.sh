mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --alpha 0.75 --hours 2 --no-cpu --no-e2e --steps 2 --warmup 1 > gpurun_out/bench_c3_lts_2h.json 2> gpurun_out/bench_c3_lts_2h.err
timeout 300 python bench.py --alpha 0.75 --hours 2 --no-cpu --no-e2e --steps 2 --warmup 1 --opt 5=1 > gpurun_out/bench_c3_lts_2h_fp64rank.json 2> gpurun_out/bench_c3_lts_2h_fp64rank.err
tail -4 gpurun_out/pytest_gpu.log; for f in bench_c3_lts_2h bench_c3_lts_2h_fp64rank; do python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
print('$f', round(d['value']), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()}, d['config']['estimates_per_step'])
PY
done; tail -c 300 gpurun_out/bench_c3_lts_2h.err
