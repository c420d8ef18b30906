mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_property.py tests/test_gpu_sfold.py -x -q 2>&1 | tail -6
