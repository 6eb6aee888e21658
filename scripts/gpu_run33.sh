mkdir -p gpurun_out
timeout 700 python scripts/parity_count.py 24 c3 > gpurun_out/parity_count_c3_24h_lazy.json 2> gpurun_out/parity_count.err; echo rc=$?
python - <<PY
import json
d=json.load(open('gpurun_out/parity_count_c3_24h_lazy.json')); d.pop('per_band'); print(d)
PY
tail -c 300 gpurun_out/parity_count.err
