python bench.py > gpurun_out/bench_latest.json 2> gpurun_out/bench_latest.err || { tail -20 gpurun_out/bench_latest.err; exit 1; }
python -c "
import json; d=json.load(open('gpurun_out/bench_latest.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['frame_tail']['ms_per_step'], d['prefill']['ms'], d['frame_loop']['ms_per_frame_sync_free'], d['frame_loop']['ms_per_frame_host_indexed'])"
