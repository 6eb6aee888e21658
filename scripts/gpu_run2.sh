mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python scripts/tc_experiment.py 1048576 3 > gpurun_out/tc_experiment_r3.json 2> gpurun_out/tc_experiment.err
timeout 300 python scripts/tc_experiment.py 1048576 9 > gpurun_out/tc_experiment_r9.json 2>> gpurun_out/tc_experiment.err
for k in tc_dft_kernel scalar_dft_kernel; do
  timeout 300 ncu --set full --clock-control none --import-source on -k "regex:$k" -s 2 -c 2 -f -o gpurun_out/r2_$k \
     python scripts/tc_experiment.py 262144 3 > gpurun_out/ncu_$k.log 2>&1
  ncu -i gpurun_out/r2_$k.ncu-rep --page raw --csv > gpurun_out/r2_$k.raw.csv 2>/dev/null
  python scripts/ncu_summary.py gpurun_out/r2_$k.raw.csv > gpurun_out/r2_$k.ncu_summary.txt 2>&1
done
timeout 200 python bench.py --workload c1 > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err
timeout 400 python bench.py --workload c4 --hours 72 --steps 1 --warmup 1 --e2e-in-flight 1 > gpurun_out/bench_c4_72h.json 2> gpurun_out/bench_c4_72h.err
timeout 400 python bench.py --workload c5 --narrays 8 --steps 1 --warmup 1 > gpurun_out/bench_c5_8.json 2> gpurun_out/bench_c5_8.err
timeout 900 python bench.py --workload c4 --steps 1 --warmup 1 --e2e-in-flight 1 --no-cpu > gpurun_out/bench_c4_720h.json 2> gpurun_out/bench_c4_720h.err
timeout 600 python bench.py --workload c5 --steps 1 --warmup 1 --no-cpu > gpurun_out/bench_c5_64.json 2> gpurun_out/bench_c5_64.err
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/tc_experiment_r3.json; tail -c 300 gpurun_out/*.err
