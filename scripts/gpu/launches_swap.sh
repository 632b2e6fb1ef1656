for o in 0 1; do echo "== gemm_swap=$o"; python scripts/profile_step.py --part dec --steps 3 --launches --opt gemm_swap=$o 2>&1 | grep -E "M=64 |ms/step" | head -24; done
