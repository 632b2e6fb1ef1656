timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err || tail -5 gpurun_out/bench_final.err
python -c "
import json; d=json.load(open('gpurun_out/bench_final.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'], d['frame_tail']['ms_per_step'], d['prefill']['ms'], d['frame_loop']['ms_per_frame_sync_free'], d['frame_loop']['ms_per_frame_host_indexed'], d['clocks'])"
python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | head -c 200
