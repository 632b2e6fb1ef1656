# round 2: ncu --set full of both k_chain launches of one frame-step (head solver loop + decoder stage 0) + launch list
mkdir -p gpurun_out
python scripts/profile_step.py --steps 1 > gpurun_out/r2_step_plain.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_step.csv python scripts/profile_step.py --steps 1 > gpurun_out/r2_ncu_launch_step.log 2>&1
ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:k_chain -c 2 -f -o gpurun_out/r2_prof_chain python scripts/profile_step.py --steps 1 > gpurun_out/r2_ncu_chain.log 2>&1
cat gpurun_out/r2_step_plain.log; ls -la gpurun_out/r2_prof_chain.ncu-rep
