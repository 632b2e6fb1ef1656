python bench.py > gpurun_out/bench_v11.json 2> gpurun_out/bench_v11.err || { tail -20 gpurun_out/bench_v11.err; exit 1; }
python -c "
import json; d=json.load(open('gpurun_out/bench_v11.json')); print(d['value'], d['ms_per_step'], d['e2e'], d['frame_tail'], d['prefill'])"
