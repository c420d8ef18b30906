mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r02_gpu_tests.log
timeout 300 python tools/exp_kernels.py rs104 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"rs_x2_crc_fast" -s 2 -c 1 -o gpurun_out/r02_x2fast_prof python tools/exp_kernels.py rs104 > gpurun_out/ncu2.log 2>&1
cat gpurun_out/plain.log
