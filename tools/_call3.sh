mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sfold.py tests/test_gpu_parity.py -x -q 2>&1 | tail -5
timeout 300 python tools/exp_kernels.py fused 128 2>&1 | tee gpurun_out/r02_fused_exp.log
GPU_EC_X2_OLD=1 timeout 300 python tools/exp_kernels.py fused 128 2>&1 | tail -1
