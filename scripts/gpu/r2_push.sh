# round 2: push grid barrier in k_chain - parity, then head / decoder / frame-step with and without it
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_head_gpu.py tests/test_golden.py tests/test_decoder_gpu.py -x -q -m gpu 2>&1 | tail -3
for o in 1 0 1 0; do
  timeout 300 python scripts/profile_step.py --part head --steps 40 --opt chain_push_barrier=$o 2>&1 | tail -1
done
for o in 1 0; do
  timeout 300 python scripts/profile_step.py --part dec --steps 40 --opt chain_push_barrier=$o 2>&1 | tail -1
  timeout 300 python scripts/profile_step.py --steps 40 --opt chain_push_barrier=$o 2>&1 | tail -1
done
