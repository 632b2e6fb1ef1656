timeout 600 python -m pytest tests/test_frameloop_gpu.py -x -q -m gpu 2>&1 | tail -30
