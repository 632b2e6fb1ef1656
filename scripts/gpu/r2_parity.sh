# round 2: parity / robustness additions first (verbose), then the whole GPU suite
timeout 900 python -m pytest tests/test_full_size_golden.py tests/test_head_gpu.py tests/test_dropin.py -x -q -s -m gpu -k "full_size or bitwise or program_cache or never_synchronises" 2>&1 | grep -v "^$" | tail -14
timeout 1800 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
