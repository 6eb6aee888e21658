mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
free -g | head -2 > gpurun_out/host_mem_$N.txt; nproc >> gpurun_out/host_mem_$N.txt
timeout 900 $TR bench.py --gpus $N --workload c4 --scaling strong --steps 2 --warmup 1 --e2e-in-flight 1 > gpurun_out/scale${N}_c4_720h_strong.json 2> gpurun_out/scale${N}_c4_720h_strong.err
timeout 600 $TR bench.py --gpus $N --steps 2 --warmup 2 > gpurun_out/scale${N}_c3_weak.json 2> gpurun_out/scale${N}_c3_weak.err
timeout 600 $TR bench.py --gpus $N --workload c5 --steps 1 --warmup 1 > gpurun_out/scale${N}_c5_64.json 2> gpurun_out/scale${N}_c5_64.err
grep -h "^{" gpurun_out/scale${N}_*.json | cut -c 1-300; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/scale${N}_*.err | tail -20
