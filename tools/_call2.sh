set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sfold.py -x -q > gpurun_out/r02_sfold_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_sfold_test.log
tail -15 gpurun_out/r02_sfold_test.log
timeout 300 python tools/exp_kernels.py sfold > gpurun_out/r02_sfold_exp.log 2>&1; cat gpurun_out/r02_sfold_exp.log
