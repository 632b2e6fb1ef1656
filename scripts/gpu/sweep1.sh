python -m pytest tests -m gpu -x -q 2>&1 | tail -1
python scripts/profile_step.py --part head --steps 50 2>&1 | tail -1
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; cat gpurun_out/bench.json
