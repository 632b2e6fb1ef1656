import math, torch
from vibevoice_b200 import _lib, synth
from vibevoice_b200.decoder import AcousticDecoderB200
def snr(a, b):
    return 10 * math.log10(b.double().pow(2).sum().item() / max((a.double() - b.double()).pow(2).sum().item(), 1e-30))
FULL = synth.DEC_FULL
W = synth.init_decoder(FULL, seed=1234)
dec = AcousticDecoderB200(W, FULL, max_rows=8)
lib = _lib.load()
g = torch.Generator().manual_seed(31)
n, T = 2, 3
lat = torch.randn(n, T, FULL.vae_dim, generator=g)
outs = {}
for chain in (0, 1):
    lib.vv_set_option(b"stage_chain", chain)
    cache = dec.new_cache(4)
    idx = torch.tensor([2, 0], device="cuda")
    outs[chain] = [dec.decode(lat[:, t:t + 1].cuda(), cache=cache, sample_indices=idx, use_cache=True).float().cpu() for t in range(T)]
for t in range(T):
    for r in range(n):
        print("frame", t, "row", r, "snr chain vs per-launch", round(snr(outs[1][t][r], outs[0][t][r]), 1), "finite", bool(torch.isfinite(outs[1][t]).all()))
from oracle import vv_oracle as O
state = O.DecoderState(4)
state_e = O.DecoderState(4)
idx = torch.tensor([2, 0])
for t in range(T):
    tru = O.decode(W, lat[:, t:t + 1], FULL.depths, FULL.ratios, FULL.vae_dim, state=state, idx=idx)
    emu = O.decode(W, lat[:, t:t + 1], FULL.depths, FULL.ratios, FULL.vae_dim, state=state_e, idx=idx, emulate_bf16=True)
    print("frame", t, "vs truth: chain", round(snr(outs[1][t], tru), 1), "per-launch", round(snr(outs[0][t], tru), 1),
          "| vs emu: chain", round(snr(outs[1][t], emu), 1), "per-launch", round(snr(outs[0][t], emu), 1), "| emu vs truth", round(snr(emu, tru), 1))
