python -m pytest tests/test_gemm_gpu.py tests/test_decoder_gpu.py tests/test_encoder_gpu.py -m gpu -x -q 2>&1 | tail -2
for cfg in "1 4" "2 2" "2 3" "2 4" "0 0"; do set -- $cfg
python scripts/profile_step.py --part dec --steps 40 --opt gemm_kps=$1 --opt gemm_stages=$2 2>&1 | tail -1
python scripts/profile_step.py --part tail --steps 40 --opt gemm_kps=$1 --opt gemm_stages=$2 2>&1 | tail -1
done
