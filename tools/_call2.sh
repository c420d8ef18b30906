set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sfold.py -x -q > gpurun_out/r02_sfold_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_sfold_test.log; tail -12 gpurun_out/r02_sfold_test.log
timeout 300 python tools/exp_kernels.py fused 128 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:rs_x2_crc_fast -s 2 -c 1 -o gpurun_out/r02_fast_prof python tools/exp_kernels.py fused 128 > gpurun_out/ncu2.log 2>&1
cat gpurun_out/plain.log
