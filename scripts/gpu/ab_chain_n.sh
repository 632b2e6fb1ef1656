for n in 1 8 32; do for o in 0 1; do timeout 300 python scripts/profile_step.py --part dec --batch $n --steps 40 --opt stage_chain=$o; done; done
for n in 1 8; do for o in 0 1; do timeout 300 python scripts/profile_step.py --part tail --batch $n --steps 40 --opt stage_chain=$o; done; done
