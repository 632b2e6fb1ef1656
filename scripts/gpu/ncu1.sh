python scripts/profile_step.py --steps 1 > gpurun_out/step_plain.log 2>&1 || exit 1
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_v9.csv python scripts/profile_step.py --steps 1 > gpurun_out/ncu_launch.log 2>&1
ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:k_chain -c 1 -f -o gpurun_out/prof_chain_v9 python scripts/profile_step.py --part head --steps 1 > gpurun_out/ncu_chain.log 2>&1
ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:k_block_fused -c 3 -f -o gpurun_out/prof_block_v9 python scripts/profile_step.py --part dec --steps 1 > gpurun_out/ncu_block.log 2>&1
cat gpurun_out/step_plain.log; ls -la gpurun_out/*.ncu-rep
