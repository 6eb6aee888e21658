mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/scale${N}_c3_weak_final.json 2> gpurun_out/scale${N}_c3_weak_final.err

grep -h "^{" gpurun_out/scale${N}_*final.json | cut -c 1-300; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/scale${N}_*final.err | tail -20
