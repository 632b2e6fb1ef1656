python bench.py > gpurun_out/bench_v10.json 2> gpurun_out/bench_v10.err || { tail -20 gpurun_out/bench_v10.err; exit 1; }
cat gpurun_out/bench_v10.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_v10_ref.json 2> gpurun_out/bench_v10_ref.err; cat gpurun_out/bench_v10_ref.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
