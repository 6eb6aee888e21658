mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_reference_arm.err
timeout 600 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
timeout 300 python bench.py --e2e-in-flight 3 --no-cpu --no-profile > gpurun_out/bench_c3_inflight3.json 2> gpurun_out/bench_c3_inflight3.err
timeout 300 python bench.py --max-lag 200 --no-cpu > gpurun_out/bench_c3_maxlag200.json 2> gpurun_out/bench_c3_maxlag200.err
timeout 300 python bench.py --workload c2 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
timeout 900 python bench.py --workload c4 --steps 1 --warmup 1 --e2e-in-flight 1 > gpurun_out/bench_c4_720h.json 2> gpurun_out/bench_c4_720h.err
timeout 600 python bench.py --workload c5 --steps 1 --warmup 1 > gpurun_out/bench_c5_64.json 2> gpurun_out/bench_c5_64.err
tail -3 gpurun_out/pytest_gpu.log; grep -h "^{" gpurun_out/bench_c3.json | cut -c 1-300; tail -c 300 gpurun_out/*.err
