timeout 600 python -m pytest tests/test_decoder_gpu.py tests/test_encoder_gpu.py -x -q -m gpu 2>&1 | tail -5
python scripts/profile_step.py --part dec --steps 10 --launches --opt stage_chain=1 2>&1 | grep -E "kind=2|ms/step"
python scripts/profile_step.py --part dec --steps 10 --launches --opt stage_chain=0 2>&1 | grep -E "M=64 |k=|ms/step" | head -3
