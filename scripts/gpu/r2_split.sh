# round 2: step boundary shared by two CTAs per row pair - parity, then head timing with and without, trace
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_head_gpu.py tests/test_golden.py tests/test_full_size_golden.py tests/test_dropin.py -x -q -m gpu 2>&1 | tail -3
for o in 1 0 1 0; do
  timeout 300 python scripts/profile_step.py --part head --steps 40 --opt head_tail_split=$o 2>&1 | tail -1
done
timeout 300 python scripts/profile_step.py --steps 40 2>&1 | tail -1
export VVB200_LIB=$PWD/vibevoice_b200/libvvb200_trace.so
timeout 300 python scripts/trace_chain.py 64 > gpurun_out/r2_chain_trace_split.txt 2>&1; grep -E "tail|phase length" gpurun_out/r2_chain_trace_split.txt | cut -c1-300
