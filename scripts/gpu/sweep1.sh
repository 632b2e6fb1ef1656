for i in 1 2; do python -m pytest tests/test_head_gpu.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -1; done
python scripts/profile_step.py --part head --steps 50 2>&1 | tail -1
python scripts/profile_step.py --part head --batch 1 --steps 50 2>&1 | tail -1
