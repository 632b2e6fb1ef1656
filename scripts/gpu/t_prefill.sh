timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_golden.py tests/test_dropin.py tests/test_encoder_gpu.py -x -q -m gpu 2>&1 | tail -25
