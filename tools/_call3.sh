mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r02_gpu_tests.log
timeout 300 python tools/exp_kernels.py fused 128 2>&1
timeout 300 python tools/exp_kernels.py rs104 1024 2>&1 | head -1
