mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_no_alloc_alone.log
tail -15 gpurun_out/pytest_no_alloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
