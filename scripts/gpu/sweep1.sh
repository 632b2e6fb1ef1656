timeout 600 python -m pytest tests/test_head_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python scripts/measure_configs.py --skip 2o,5 2>&1 | grep large_head
