timeout 900 python -m pytest tests/test_decoder_gpu.py tests/test_encoder_gpu.py -x -q -m gpu -k "batch64" 2>&1 | tail -12
