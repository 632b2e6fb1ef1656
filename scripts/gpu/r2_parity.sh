# round 2: parity additions (late-joiner quirk both modes, slot guard) + whole GPU suite
timeout 600 python -m pytest tests/test_golden.py tests/test_decoder_gpu.py -x -q -m gpu -k "quirk or late_joiner or out_of_range" 2>&1 | tail -5
timeout 1800 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
