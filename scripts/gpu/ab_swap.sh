timeout 900 python -m pytest tests/test_gemm_gpu.py -x -q -m gpu -k "partial or swapped" 2>&1 | tail -6
timeout 900 python -m pytest tests/test_decoder_gpu.py tests/test_encoder_gpu.py tests/test_golden.py -x -q -m gpu 2>&1 | tail -4
for o in 0 1 0 1; do python scripts/profile_step.py --part both --steps 40 --opt gemm_swap=$o; done
for o in 0 1; do python scripts/profile_step.py --part tail --steps 40 --opt gemm_swap=$o; done
