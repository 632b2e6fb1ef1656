python -m pytest tests/test_head_gpu.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3
for sa in 4 5 8; do for sb in 4 24; do
python scripts/profile_step.py --part head --steps 20 --opt chain_stages=$sa --opt chain_wstages=$sb 2>&1 | tail -1
done; done
