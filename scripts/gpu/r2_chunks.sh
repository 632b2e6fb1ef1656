timeout 900 python -m pytest tests/test_decoder_gpu.py tests/test_encoder_gpu.py -x -q -s -m gpu 2>&1 | grep -E "passed|failed|Error|assert|bit" | tail -8
for o in pipe_chunks=1 pipe_chunks=0 pipe_chunks=1 pipe_chunks=0; do timeout 300 python scripts/profile_step.py --part dec --steps 40 --opt $o 2>&1 | tail -1; done
timeout 300 python scripts/profile_step.py --steps 40 2>&1 | tail -1
