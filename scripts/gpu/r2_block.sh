# round 2: pipelined Block1D kernel - parity, then timing with and without it
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_decoder_gpu.py -x -q -s -m gpu 2>&1 | tail -25
for o in 1 0; do
  timeout 300 python scripts/profile_step.py --part dec --steps 40 --opt block_pipe=$o 2>&1 | tail -2
done
timeout 300 python scripts/profile_step.py --steps 40 2>&1 | tail -1
