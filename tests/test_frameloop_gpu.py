"""GPU tests of the sync-free frame iteration (vibevoice_b200/frameloop.py, csrc/vv_loop.cu).

The reference (modeling_vibevoice_inference.py:586-690) compacts the rows in diffusion mode on the host and runs the
acoustic path on that subset; FrameLoopB200 runs the whole batch with inactive rows redirected to a spare cache slot.
Parity = per active row the two agree: the same frames driven through (a) FrameLoopB200.step and (b) the
index-compacted calls generate() makes (head -> decode(sample_indices) -> encode(sample_indices) -> connectors,
each already parity-tested against the oracle), on separate handles and caches.  Bookkeeping (row selection,
compaction, ring, masked scatter) is integer/byte work and must be bit-exact; the audio/embeddings agree to >= 35 dB
(the GEMM tiling depends on the row count, so not bitwise) and also against the CPU oracle run on the compacted rows.
"""
import math

import pytest
import torch

from oracle import vv_oracle as O
from vibevoice_b200 import _lib, synth
from vibevoice_b200._lib import check, ptr

pytestmark = pytest.mark.gpu

HC = synth.HeadCfg(hidden=128, layers=2)
DC = synth.DecCfg(vae_dim=HC.latent, n_filters=8, ratios=(2, 2, 2), depths=(2, 1, 1, 2))
EC = synth.EncCfg(dim=16, n_filters=8, ratios=(2, 2, 2), depths=(2, 1, 1, 2))
CA = synth.ConnCfg(input_dim=HC.latent, output_dim=HC.hidden)
CS = synth.ConnCfg(input_dim=EC.dim, output_dim=HC.hidden)
DIFF, END, OTHER = 7, 8, 3
SCALE, BIAS = 0.1962890625, -0.0703125


def _snr(got, ref):
    err = (got.double() - ref.double()).pow(2).sum().item()
    return 10 * math.log10(ref.double().pow(2).sum().item() / max(err, 1e-30))


def _handles(n):
    from vibevoice_b200.decoder import AcousticDecoderB200
    from vibevoice_b200.encoder import SemanticEncoderB200, SpeechConnectorB200
    from vibevoice_b200.head import DiffusionHeadB200

    W = dict(head=synth.init_head(HC, seed=5), dec=synth.init_decoder(DC, seed=5), enc=synth.init_encoder(EC, seed=5),
             ca=synth.init_connector(CA, seed=5, tag="a"), cs=synth.init_connector(CS, seed=5, tag="s"))
    head = DiffusionHeadB200(W["head"], HC, max_rows=2 * n)
    dec = AcousticDecoderB200(W["dec"], DC, max_rows=max(n, 8))
    enc = SemanticEncoderB200(W["enc"], EC, max_rows=max(n, 8))
    ca = SpeechConnectorB200(W["ca"], CA, max_rows=max(n, 8))
    cs = SpeechConnectorB200(W["cs"], CS, max_rows=max(n, 8))
    return W, head, dec, enc, ca, cs


def test_frame_select_matches_torch_bitwise():
    lib = _lib.load()
    g = torch.Generator().manual_seed(0)
    for n in (1, 5, 64, 1024, 2500):
        tok = torch.randint(6, 9, (n,), generator=g).cuda()
        fin = (torch.rand(n, generator=g) < 0.3).cuda()
        slots = torch.empty(n, dtype=torch.int64, device="cuda")
        act = torch.empty(n, dtype=torch.bool, device="cuda")
        comp = torch.empty(n, dtype=torch.int64, device="cuda")
        cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
        check(lib.vv_frame_select(ptr(tok), ptr(fin), n, DIFF, n, ptr(slots), ptr(act), ptr(comp), ptr(cnt), None), "sel")
        torch.cuda.synchronize()
        want = ~fin & (tok == DIFF)
        idx = torch.arange(n, device="cuda")[want]                       # the reference's expression (:588)
        assert torch.equal(act, want)
        assert torch.equal(slots, torch.where(want, torch.arange(n, device="cuda"), torch.full_like(slots, n)))
        assert int(cnt) == idx.numel()
        assert torch.equal(comp[:idx.numel()], idx) and bool((comp[idx.numel():] == -1).all())
    # either input alone
    check(lib.vv_frame_select(ptr(tok), None, n, DIFF, n, None, ptr(act), None, None, None), "sel")
    torch.cuda.synchronize()
    assert torch.equal(act, tok == DIFF)
    check(lib.vv_frame_select(None, ptr(fin), n, 0, n, None, ptr(act), None, None, None), "sel")
    torch.cuda.synchronize()
    assert torch.equal(act, ~fin)


@pytest.mark.parametrize("dtype,hop", [(torch.bfloat16, 3200), (torch.float32, 8), (torch.bfloat16, 5)])
def test_ring_append_and_embed_scatter_bitwise(dtype, hop):
    lib = _lib.load()
    io = _lib.VV_DTYPE_BF16 if dtype == torch.bfloat16 else _lib.VV_DTYPE_F32
    n, cap, F = 6, 4, 6
    g = torch.Generator().manual_seed(1)
    ring = torch.zeros(n, cap, hop, dtype=dtype, device="cuda")
    pos = torch.zeros(n, dtype=torch.int32, device="cuda")
    want = [[] for _ in range(n)]
    for f in range(F):
        chunk = torch.randn(n, 1, hop, generator=g).to(dtype).cuda()
        act = (torch.rand(n, generator=g) < 0.6)
        act[0] = True                                                    # row 0 overflows the ring (6 > cap)
        check(lib.vv_ring_append(ptr(chunk), ptr(act.cuda()), n, hop, io, ptr(ring), ptr(pos), cap, None), "ring")
        for i in range(n):
            if act[i]:
                want[i].append(chunk[i, 0].cpu())
    torch.cuda.synchronize()
    assert pos.cpu().tolist() == [len(w) for w in want]
    for i in range(n):
        k = min(len(want[i]), cap)
        if k:
            assert torch.equal(ring[i, :k].cpu(), torch.stack(want[i][:k]))
        assert bool((ring[i, k:] == 0).all())
    # masked embedding update
    H = 96
    emb = torch.randn(n, H, generator=g).bfloat16().cuda()
    for ddt, dio in ((torch.bfloat16, _lib.VV_DTYPE_BF16), (torch.float32, _lib.VV_DTYPE_F32)):
        dst0 = torch.randn(n, 1, H, generator=g).to(ddt).cuda()
        dst = dst0.clone()
        act = torch.tensor([True, False, True, True, False, False]).cuda()
        check(lib.vv_embed_scatter(ptr(emb), ptr(act), n, H, ptr(dst), dio, None), "scatter")
        torch.cuda.synchronize()
        ref = dst0.clone()
        ref[act] = emb.to(ddt).unsqueeze(1)[act]                         # next_inputs_embeds[idx] = diffusion_embeds (:687)
        assert torch.equal(dst, ref)


def test_frame_loop_equals_index_compacted_path():
    from vibevoice_b200.frameloop import FrameLoopB200, PinnedChunkStream

    n, F = 5, 7
    W, head, dec, enc, ca, cs = _handles(n)
    loop = FrameLoopB200(head, dec, enc, ca, cs, batch=n, max_frames=F, speech_diffusion_id=DIFF,
                         scaling_factor=SCALE, bias_factor=BIAS, cfg_scale=1.3, num_steps=4)
    stream = PinnedChunkStream(n, DC.hop)
    # (b) the reference's call sequence on its own handles and caches
    W2, head2, dec2, enc2, ca2, cs2 = _handles(n)
    head2.set_schedule(4, "dpmsolver++")
    acache, scache = dec2.new_cache(n), enc2.new_cache(n)
    # (c) CPU oracle state
    dstate, estate = O.DecoderState(n), O.DecoderState(n)

    g = torch.Generator().manual_seed(3)
    tokens = [[DIFF, DIFF, DIFF, OTHER, DIFF], [DIFF, OTHER, DIFF, DIFF, DIFF], [DIFF, DIFF, END, DIFF, DIFF],
              [DIFF, DIFF, DIFF, DIFF, DIFF], [OTHER, OTHER, OTHER, OTHER, OTHER], [DIFF, DIFF, DIFF, DIFF, DIFF],
              [DIFF, END, DIFF, DIFF, OTHER]]
    finished = torch.zeros(n, dtype=torch.bool)
    chunks_b = [[] for _ in range(n)]
    got_stream = [[] for _ in range(n)]
    for f in range(F):
        tok = torch.tensor(tokens[f])
        if f == 3:
            finished[4] = True                                           # row 4 is done: DIFF tokens no longer count
        end = tok == END
        pos_h = torch.randn(n, HC.hidden, generator=g).bfloat16()
        neg_h = torch.randn(n, HC.hidden, generator=g).bfloat16()
        noise = torch.randn(2 * n, HC.latent, generator=g)
        emb0 = torch.randn(n, 1, HC.hidden, generator=g).bfloat16()
        # ---- (a)
        loop.end_speech(end.cuda())
        nie = emb0.clone().cuda()
        chunk, active = loop.step(tok.cuda(), finished.cuda(), pos_h.cuda(), neg_h.cuda(), nie, noise=noise.cuda())
        stream.put(chunk, active)
        hc, ha = stream.get()
        for i in range(n):
            if ha[i]:
                got_stream[i].append(hc[i].clone())
        # ---- (b)
        want_active = ~finished & (tok == DIFF)
        assert torch.equal(active.cpu(), want_active)
        if end.any():
            acache.set_to_zero(end.nonzero().squeeze(1))
            scache.set_to_zero(end.nonzero().squeeze(1))
            dstate.set_to_zero(end.nonzero().squeeze(1))
            estate.set_to_zero(end.nonzero().squeeze(1))
        idx = torch.arange(n)[want_active]
        nie_cpu = nie.cpu()
        assert torch.equal(nie_cpu[~want_active], emb0[~want_active])      # untouched rows: bit-identical
        if idx.numel() == 0:
            continue
        k = idx.numel()
        nz = torch.cat([noise[idx], noise[idx]])                         # only the first half is used (:716-719)
        lat = head2.sample_speech_tokens(pos_h[idx].cuda(), neg_h[idx].cuda(), 1.3, noise=nz).unsqueeze(1)
        ch = dec2.decode(lat, cache=acache, sample_indices=idx.cuda(), use_cache=True, scale=SCALE, bias=BIAS)
        feat = enc2.encode(ch, cache=scache, sample_indices=idx.cuda(), use_cache=True).mean
        emb = ca2(lat) + cs2(feat)
        for j, i in enumerate(idx.tolist()):
            chunks_b[i].append(ch[j, 0].float().cpu())
        assert _snr(loop.last_latent.float().cpu()[idx], lat.float().cpu()) >= 35.0
        assert _snr(chunk.float().cpu()[idx], ch.float().cpu()) >= 35.0
        assert _snr(nie_cpu.float()[idx], emb.float().cpu()) >= 35.0
        # ---- (c) the CPU oracle on the compacted rows, from the latents of path (a)
        la = loop.last_latent.float().cpu()[idx]
        o_ch = O.decode(W["dec"], la / SCALE - BIAS, DC.depths, DC.ratios, DC.vae_dim, state=dstate, idx=idx)
        o_feat = O.encode(W["enc"], o_ch, EC.depths, EC.ratios, EC.eps, state=estate, idx=idx)
        o_emb = O.speech_connector(W["ca"], la, CA.eps) + O.speech_connector(W["cs"], o_feat, CS.eps)
        assert _snr(chunk.float().cpu()[idx], o_ch) >= 28.0, (f, _snr(chunk.float().cpu()[idx], o_ch))
        assert _snr(nie_cpu.float()[idx], o_emb) >= 28.0
    waves = loop.audio()
    counts = [sum(1 for f in range(F) if tokens[f][i] == DIFF and not (i == 4 and f >= 3)) for i in range(n)]
    assert loop.pos.cpu().tolist() == counts
    for i in range(n):
        if counts[i] == 0:
            assert waves[i] is None
            continue
        assert waves[i].shape == (1, counts[i] * DC.hop)
        assert torch.equal(waves[i].cpu().reshape(-1), torch.cat(got_stream[i]).reshape(-1))   # ring == streamed chunks
        assert _snr(waves[i].float().cpu().reshape(-1), torch.cat(chunks_b[i])) >= 35.0
    assert int(loop.count) == sum(1 for i in range(n) if tokens[F - 1][i] == DIFF and not finished[i])
    loop.reset()
    assert loop.pos.sum().item() == 0


def test_ring_overflow_is_reported():
    from vibevoice_b200.frameloop import FrameLoopB200

    n = 2
    W, head, dec, enc, ca, cs = _handles(n)
    loop = FrameLoopB200(head, dec, enc, ca, cs, batch=n, max_frames=1, speech_diffusion_id=DIFF, num_steps=2)
    tok = torch.full((n,), DIFF).cuda()
    h = torch.randn(n, HC.hidden).bfloat16().cuda()
    for _ in range(2):
        loop.step(tok, None, h, h)
    with pytest.raises(_lib.VVError):
        loop.audio()
