python -m pytest tests/test_dropin.py -m gpu -x -q 2>&1 | tail -3
python scripts/profile_step.py --part tail --steps 30 2>&1 | tail -1
python scripts/profile_step.py --part all --steps 30 2>&1 | tail -1
python scripts/profile_step.py --part both --steps 30 2>&1 | tail -1
