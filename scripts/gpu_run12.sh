mkdir -p gpurun_out
timeout 900 python scripts/parity_count.py 1.0 c3 > gpurun_out/parity_count_c3_1h.json 2> gpurun_out/parity_count_c3_1h.err
timeout 600 python scripts/parity_count.py 2.0 c2 > gpurun_out/parity_count_c2_2h.json 2> gpurun_out/parity_count_c2_2h.err
python - <<PY
import json
for f in ('parity_count_c3_1h','parity_count_c2_2h'):
    try:
        d=json.loads([l for l in open('gpurun_out/'+f+'.json') if l.startswith('{')][-1])
        print(f, {k:v for k,v in d.items() if k!='per_band'})
    except Exception as e: print(f,'ERR',e)
PY
tail -c 500 gpurun_out/parity_count*.err
