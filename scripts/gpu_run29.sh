# multi-GPU check of the final build: N = $1
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514"
timeout 600 $TR bench.py --gpus $N --steps 4 --warmup 3 > gpurun_out/scale${N}_c3_weak.json 2> gpurun_out/scale${N}_c3_weak.err
timeout 600 $TR bench.py --gpus $N --workload c4 --scaling strong --hours ${2:-72} --steps 1 --warmup 1 --no-cpu > gpurun_out/scale${N}_c4_strong.json 2> gpurun_out/scale${N}_c4_strong.err
for f in scale${N}_c3_weak scale${N}_c4_strong; do python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
    print('$f', round(d['value']), 'e2e', (d.get('e2e') or {}).get('value'), round(d['ms_per_step'],2), d.get('multi_gpu_check'))
except Exception as e:
    print('$f FAILED', e); print(open('gpurun_out/$f.err').read()[-800:])
PY
done
