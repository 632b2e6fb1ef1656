timeout 600 python -m pytest tests/test_decoder_gpu.py -x -q -m gpu -k "stage_chain or full_decoder" 2>&1 | tail -8 || exit 1
for o in 0 1 0 1; do timeout 300 python scripts/profile_step.py --part both --steps 40 --opt stage_chain=$o; done
for o in 0 1; do timeout 300 python scripts/profile_step.py --part head --steps 40 --opt stage_chain=$o; done
for o in 0 1; do timeout 300 python scripts/profile_step.py --part tail --steps 40 --opt stage_chain=$o; done
