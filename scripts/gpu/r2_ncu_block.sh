# round 2: launch list of the decoder step + ncu --set full of k_block_pipe (3 launches: C = 128 / 64 / 32 blocks of a stage each)
mkdir -p gpurun_out
python scripts/profile_step.py --part dec --steps 1 > gpurun_out/r2_dec_plain.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_dec.csv python scripts/profile_step.py --part dec --steps 1 > gpurun_out/r2_ncu_launch.log 2>&1
ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:k_block_pipe -c 9 -f -o gpurun_out/r2_prof_block python scripts/profile_step.py --part dec --steps 1 > gpurun_out/r2_ncu_block.log 2>&1
cat gpurun_out/r2_dec_plain.log; ls -la gpurun_out/*.ncu-rep
