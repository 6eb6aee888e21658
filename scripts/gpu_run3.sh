mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
timeout 300 python bench.py --single-stage --no-cpu > gpurun_out/bench_c3_single_stage.json 2> gpurun_out/bench_c3_single_stage.err
timeout 300 python bench.py --noise-only --no-cpu > gpurun_out/bench_c3_noise_only.json 2> gpurun_out/bench_c3_noise_only.err
timeout 300 python bench.py --workload c2 --no-cpu > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
timeout 400 python bench.py --workload c4 --hours 72 --steps 1 --warmup 1 --e2e-in-flight 1 --no-cpu > gpurun_out/bench_c4_72h.json 2> gpurun_out/bench_c4_72h.err
timeout 900 bash scripts/ncu_traffic.sh c3 2 24 > gpurun_out/ncu_traffic.out 2>&1
timeout 400 bash scripts/ncu_kernel.sh xf_pair_kernel r2_xf_pair 12 --hours 1 > gpurun_out/ncu_pair.out 2>&1
timeout 400 bash scripts/ncu_kernel.sh xf_verify_kernel r2_xf_verify 6 --hours 1 > gpurun_out/ncu_verify.out 2>&1
timeout 400 bash scripts/ncu_kernel.sh lts_kernel r2_lts 0 --workload c4 --hours 6 > gpurun_out/ncu_lts.out 2>&1
cp profiles/*.ncu_summary.txt gpurun_out/ 2>/dev/null
tail -3 gpurun_out/pytest_gpu.log; head -c 600 gpurun_out/bench_c3.json; tail -c 300 gpurun_out/*.err
