python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python scripts/profile_step.py --part head --steps 20 2>&1 | tail -1
python scripts/profile_step.py --part dec --steps 20 2>&1 | tail -1
python scripts/profile_step.py --part both --steps 20 2>&1 | tail -1
