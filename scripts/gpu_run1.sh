mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/gpu.txt; nproc >> gpurun_out/gpu.txt; free -g >> gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "rc=$?" >> gpurun_out/bench_c3.err
timeout 300 python bench.py --max-lag 200 --no-cpu > gpurun_out/bench_c3_maxlag200.json 2> gpurun_out/bench_c3_maxlag200.err
timeout 300 python bench.py --workload c2 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
timeout 600 bash scripts/ncu_traffic.sh c3 2 24 > gpurun_out/ncu_traffic.out 2>&1
timeout 600 bash scripts/sanitize.sh > gpurun_out/sanitize.out 2>&1
tail -5 gpurun_out/pytest_gpu.log; tail -c 600 gpurun_out/bench_c3.err; head -c 1500 gpurun_out/bench_c3.json
