mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sfold.py tests/test_gpu_parity.py -x -q 2>&1 | tail -5
timeout 300 python tools/exp_kernels.py fused63 128 2>&1 | tee gpurun_out/r02_fused63_exp.log
timeout 300 python tools/exp_kernels.py fused63 128 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:rs_lut8_crc -s 2 -c 1 -o gpurun_out/r02_fused63_prof python tools/exp_kernels.py fused63 128 > gpurun_out/ncu2.log 2>&1
