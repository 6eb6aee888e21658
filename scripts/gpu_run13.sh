mkdir -p gpurun_out
timeout 2400 python scripts/parity_count.py 24.0 c3 > gpurun_out/parity_count_c3_24h.json 2> gpurun_out/parity_count_c3_24h.err
timeout 900 python scripts/parity_count.py 72.0 c4 > gpurun_out/parity_count_c4_72h.json 2> gpurun_out/parity_count_c4_72h.err
python - <<PY
import json
for f in ('parity_count_c3_24h','parity_count_c4_72h'):
    try:
        d=json.loads([l for l in open('gpurun_out/'+f+'.json') if l.startswith('{')][-1])
        print(f, {k:v for k,v in d.items() if k!='per_band'})
    except Exception as e: print(f,'ERR',e)
PY
tail -c 500 gpurun_out/parity_count*.err
