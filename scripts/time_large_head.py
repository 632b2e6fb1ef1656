import sys, os, torch
sys.path.insert(0, os.getcwd())
from vibevoice_b200 import _lib, synth
from vibevoice_b200.head import DiffusionHeadB200
lib = _lib.load()
cfg = synth.HEAD_LARGE
W = synth.init_head(cfg, seed=7)
for n in (1, 16, 64):
    for chain in (0, 1):
        lib.vv_set_option(b"head_chain", chain)
        head = DiffusionHeadB200(W, cfg, max_rows=2 * n)
        cond, neg, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent, seed=3)
        cond, neg = cond.cuda().bfloat16(), neg.cuda().bfloat16()
        for _ in range(3): head.sample_speech_tokens(cond, neg, 1.3, num_steps=10, noise=noise)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): head.sample_speech_tokens(cond, neg, 1.3, num_steps=10, noise=noise)
        e1.record(); torch.cuda.synchronize()
        print(f"large head n={n} head_chain={chain}: {e0.elapsed_time(e1) / 10:.3f} ms")
        del head
