# round 2: two MMA issuers in k_chain + program table out of shared memory - parity, then head / decoder / frame-step timing, trace
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_head_gpu.py tests/test_golden.py tests/test_decoder_gpu.py tests/test_encoder_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python scripts/profile_step.py --part head --steps 40 2>&1 | tail -1
timeout 300 python scripts/profile_step.py --part dec --steps 40 2>&1 | tail -1
timeout 300 python scripts/profile_step.py --steps 40 2>&1 | tail -1
export VVB200_LIB=$PWD/vibevoice_b200/libvvb200_trace.so
timeout 300 python scripts/trace_chain.py 64 > gpurun_out/r2_chain_trace_v5.txt 2>&1; tail -22 gpurun_out/r2_chain_trace_v5.txt | cut -c1-200
