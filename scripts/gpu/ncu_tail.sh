python scripts/profile_step.py --part tail --steps 20 > gpurun_out/tail_plain.log 2>&1 || { tail gpurun_out/tail_plain.log; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_tail.csv python scripts/profile_step.py --part tail --steps 1 > gpurun_out/ncu_tail.log 2>&1
cat gpurun_out/tail_plain.log | tail -3
