timeout 900 python -m pytest tests/test_encoder_gpu.py tests/test_prefill_gpu.py tests/test_golden.py tests/test_frameloop_gpu.py tests/test_dropin.py -x -q -m gpu 2>&1 | tail -4
python scripts/profile_step.py --part tail --steps 40 --opt enc_splitk=0
python scripts/profile_step.py --part tail --steps 40 --opt enc_splitk=1
python scripts/profile_step.py --part tail --steps 40 --opt enc_splitk=0
python scripts/profile_step.py --part tail --steps 40 --opt enc_splitk=1
