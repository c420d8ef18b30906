mkdir -p gpurun_out
timeout 600 python tools/exp_kernels.py bigcall 2>&1 | tail -1 | tee gpurun_out/r02_bigcall.log
GPU_EC_NO_CHUNKED_COPY=1 timeout 600 python tools/exp_kernels.py bigcall 2>&1 | tail -1 | tee -a gpurun_out/r02_bigcall.log
