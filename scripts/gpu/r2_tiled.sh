# round 2: tiled stage program - parity, then A/B timing
timeout 900 python -m pytest tests/test_decoder_gpu.py tests/test_encoder_gpu.py -x -q -s -m gpu 2>&1 | grep -E "stage_tiled|passed|failed|Error|assert" | tail -12
for o in stage_tiled=4096 stage_tiled=0; do
  timeout 300 python scripts/profile_step.py --part dec --steps 40 --opt $o 2>&1 | tail -1
done
timeout 300 python scripts/profile_step.py --part tail --steps 40 2>&1 | tail -1
timeout 300 python scripts/profile_step.py --part tail --steps 40 --opt stage_tiled=0 2>&1 | tail -1
timeout 300 python scripts/profile_step.py --steps 40 2>&1 | tail -1
