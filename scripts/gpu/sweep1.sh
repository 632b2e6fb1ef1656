python -m pytest tests/test_encoder_gpu.py tests/test_golden.py tests/test_abi.py -x -q 2>&1 | tail -25
