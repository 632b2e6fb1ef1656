"""N>1 host path on CPU: world_size-2 gloo processes shard the rows, run the frame loop with the
oracle standing in for the CUDA operators (tests may use it as the engine: it is the checker), and
rank 0's gathered waveforms must equal the single-process result (fp32 round-off only: BLAS blocks
differently for different batch sizes)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import vv_oracle as O
from vibevoice_b200 import shard, synth

HC = synth.HeadCfg(hidden=64, layers=1)
DC = synth.DEC_TINY
N_TOTAL, FRAMES, STEPS = 5, 3, 3      # 5 rows over 2 ranks: uneven shards (3 + 2)


def _ops(n_slots):
    Wh = synth.init_head(HC, seed=11)
    Wd = synth.init_decoder(synth.DecCfg(vae_dim=HC.latent, n_filters=DC.n_filters, ratios=DC.ratios, depths=DC.depths),
                            seed=11)
    state = O.DecoderState(max(n_slots, 1))

    def sample(c, ng, cfg, nz):
        return O.sample_speech_tokens(Wh, c, ng, cfg, nz, steps=STEPS, t_dtype=torch.float32, layers=HC.layers)

    def decode_step(lat, rows):
        return O.decode(Wd, lat, DC.depths, DC.ratios, HC.latent, state=state, idx=rows)

    return sample, decode_step


def _inputs():
    g = torch.Generator().manual_seed(77)
    return [(torch.randn(N_TOTAL, HC.hidden, generator=g), torch.randn(N_TOTAL, HC.hidden, generator=g),
             torch.randn(2 * N_TOTAL, HC.latent, generator=g)) for _ in range(FRAMES)]


def _run(world, rank):
    a, b = shard.row_range(N_TOTAL, world, rank)
    sample, decode_step = _ops(b - a)
    path = shard.ShardedAcousticPath(sample, decode_step, N_TOTAL, world, rank)
    for cond, neg, noise in _inputs():
        path.frame_step(cond, neg, 1.3, noise)
    return path.finish(dst=0)


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        got = _run(world, rank)
        if rank == 0:
            torch.save(got, out_path)
        else:
            assert got is None
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_row_range_partitions_everything():
    for n in (0, 1, 5, 64, 513):
        for world in (1, 2, 3, 8):
            spans = [shard.row_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.row_range(4, 2, 2)


def test_shard_noise_consumes_the_single_process_draw():
    noise = torch.arange(2 * 5 * 2, dtype=torch.float32).view(10, 2)
    parts = [shard.shard_noise(noise, 2, r) for r in range(2)]
    assert torch.equal(torch.cat([p[: p.shape[0] // 2] for p in parts]), noise[:5])
    assert torch.equal(torch.cat([p[p.shape[0] // 2:] for p in parts]), noise[5:])


def test_world2_gloo_matches_single_process(tmp_path):
    want = _run(1, 0)
    assert want.shape == (N_TOTAL, 1, DC.hop * FRAMES)
    out = str(tmp_path / "gathered.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    # same rows, same arithmetic; only the BLAS blocking differs with the batch size
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6)
