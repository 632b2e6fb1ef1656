# round 2: head (k_chain) parity + A/B timing of a vv_set_option
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_head_gpu.py tests/test_golden.py -x -q -m gpu 2>&1 | tail -3
for o in $@; do
  timeout 300 python scripts/profile_step.py --part head --steps 40 --opt $o 2>&1 | tail -1
done
timeout 300 python scripts/profile_step.py --steps 40 2>&1 | tail -1
