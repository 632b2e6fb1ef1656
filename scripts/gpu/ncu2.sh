python scripts/profile_step.py --steps 1 > gpurun_out/step_plain13.log 2>&1 || exit 1
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_v13.csv python scripts/profile_step.py --steps 1 > gpurun_out/ncu_launch13.log 2>&1
ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:k_chain -c 1 -f -o gpurun_out/prof_stage_chain_v13 python scripts/profile_step.py --part dec --steps 1 > gpurun_out/ncu_stage13.log 2>&1
cat gpurun_out/step_plain13.log; ls -la gpurun_out/*v13*
