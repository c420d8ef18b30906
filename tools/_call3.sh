timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -k "reconstruct or decode or encode_config1" 2>&1 | tail -3
