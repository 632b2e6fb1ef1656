timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 120 python scripts/profile_step.py --part both --steps 40 2>&1 | tail -1
