# round 2: bulk-staged mixer kernel - parity, then decoder step with and without it
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_decoder_gpu.py tests/test_encoder_gpu.py tests/test_golden.py tests/test_full_size_golden.py -x -q -m gpu 2>&1 | tail -3
for o in 1 0 1 0; do
  timeout 300 python scripts/profile_step.py --part dec --steps 40 --opt dw_bulk=$o 2>&1 | tail -1
done
timeout 300 python scripts/profile_step.py --part dec --steps 3 --log > gpurun_out/r2_launch_log_dec3.txt 2>&1
grep -E "k_block_dw|total" gpurun_out/r2_launch_log_dec3.txt | cut -c1-100 | tail -12
