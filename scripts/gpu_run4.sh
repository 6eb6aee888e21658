mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 2 --warmup 2 > gpurun_out/scale2_c3_weak.json 2> gpurun_out/scale2_c3_weak.err
timeout 600 $TR bench.py --gpus 2 --workload c4 --hours 72 --scaling strong --steps 1 --warmup 1 --e2e-in-flight 1 > gpurun_out/scale2_c4_72h_strong.json 2> gpurun_out/scale2_c4_72h_strong.err
timeout 600 $TR bench.py --gpus 2 --workload c5 --narrays 8 --steps 1 --warmup 1 > gpurun_out/scale2_c5_8.json 2> gpurun_out/scale2_c5_8.err
timeout 300 python scripts/screen_audit.py > gpurun_out/screen_audit.json 2> gpurun_out/screen_audit.err
timeout 300 ncu --set full --clock-control none --kernel-name-base demangled -k "regex:tc_dft_kernel<2>" -s 2 -c 1 -f -o gpurun_out/r2_tc_dft_split python scripts/tc_experiment.py 262144 3 > gpurun_out/ncu_tc2.log 2>&1
ncu -i gpurun_out/r2_tc_dft_split.ncu-rep --page raw --csv > gpurun_out/r2_tc_dft_split.raw.csv 2>/dev/null; python scripts/ncu_summary.py gpurun_out/r2_tc_dft_split.raw.csv > gpurun_out/r2_tc_dft_kernel_split.ncu_summary.txt 2>&1
head -c 400 gpurun_out/scale2_c3_weak.json; tail -c 400 gpurun_out/*.err
