# round 2: pipelined Block1D kernel - parity, timing with and without it, per-tile trace
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_decoder_gpu.py -x -q -m gpu 2>&1 | tail -3
for o in 1 0; do
  timeout 300 python scripts/profile_step.py --part dec --steps 40 --opt block_pipe=$o 2>&1 | tail -1
done
timeout 300 python scripts/profile_step.py --steps 40 2>&1 | tail -1
export VVB200_LIB=$PWD/vibevoice_b200/libvvb200_trace.so
for s in 7 6; do timeout 200 python scripts/trace_pipe.py 64 $s 2>&1 | grep -E "last-stage|cta 77|Error|error" | cut -c1-220; done
