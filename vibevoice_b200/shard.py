"""Data-parallel sharding of the acoustic path: one process per GPU, utterances (batch rows)
partitioned contiguously, weights replicated, no collective in the frame loop, one final gather.

Rows are independent everywhere on the path (RMSNorm is per row / per time step, CFG pairs row i
with row n+i of the same utterance, the conv cache is per sample_idx — SURVEY.md §8(e)), so the
only communication is collecting the finished waveforms on rank 0.  Works with any
torch.distributed backend (nccl on the B200 box, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist


def row_range(n_total: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous [start, stop) of the rows rank owns; the remainder goes to the lowest ranks."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(n_total, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def shard_rows(t: torch.Tensor, world: int, rank: int) -> torch.Tensor:
    a, b = row_range(t.shape[0], world, rank)
    return t[a:b]


def shard_noise(noise: torch.Tensor, world: int, rank: int) -> torch.Tensor:
    """noise is the reference's single torch.randn(2n, L) draw (modeling_vibevoice_inference.py:716):
    rows [0,n) seed the latents, rows [n,2n) are carried but never read.  Each rank gets the rows of
    its utterances from both halves, so a sharded run consumes exactly the single-process draw."""
    n = noise.shape[0] // 2
    a, b = row_range(n, world, rank)
    return torch.cat([noise[a:b], noise[n + a:n + b]], dim=0)


def gather_rows(local: torch.Tensor, n_total: int, dst: int = 0, group=None) -> Optional[torch.Tensor]:
    """The final gather: every rank contributes its [n_local, ...] block; dst returns [n_total, ...].
    Uneven shards are padded to the largest block (gather needs equal sizes) and trimmed."""
    if not dist.is_available() or not dist.is_initialized():
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if world == 1:
        return local
    sizes = [row_range(n_total, world, r) for r in range(world)]
    mx = max(b - a for a, b in sizes)
    pad = local
    if local.shape[0] < mx:
        pad = torch.cat([local, local.new_zeros((mx - local.shape[0],) + tuple(local.shape[1:]))], dim=0)
    pad = pad.contiguous()
    bufs: Optional[List[torch.Tensor]] = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst, group=group)
    if rank != dst:
        return None
    return torch.cat([bufs[r][: sizes[r][1] - sizes[r][0]] for r in range(world)], dim=0)


class ShardedAcousticPath:
    """Runs the per-frame path on this rank's rows.

    ``sample(cond, neg, cfg, noise) -> latent`` and ``decode_step(latent, rows) -> wave`` are the two
    operators of the path (on the GPU: DiffusionHeadB200.sample_speech_tokens and
    AcousticDecoderB200.decode with the streaming cache); they are injected so the host logic can be
    exercised without a GPU.
    """

    def __init__(self, sample: Callable, decode_step: Callable, n_total: int, world: int = 1, rank: int = 0):
        self.sample, self.decode_step = sample, decode_step
        self.n_total, self.world, self.rank = n_total, world, rank
        self.a, self.b = row_range(n_total, world, rank)
        self.chunks: List[torch.Tensor] = []

    @property
    def n_local(self) -> int:
        return self.b - self.a

    def frame_step(self, cond: torch.Tensor, neg: torch.Tensor, cfg_scale: float, noise: torch.Tensor) -> None:
        """cond/neg [n_total,H], noise [2*n_total,L] are the GLOBAL per-frame tensors; only this rank's
        rows are touched."""
        if self.n_local == 0:
            return
        c, ng = cond[self.a:self.b], neg[self.a:self.b]
        nz = shard_noise(noise, self.world, self.rank)
        lat = self.sample(c, ng, cfg_scale, nz)
        rows = torch.arange(self.n_local, device=lat.device)      # cache slots are rank-local
        self.chunks.append(self.decode_step(lat.unsqueeze(1), rows))

    def finish(self, dst: int = 0) -> Optional[torch.Tensor]:
        """Concatenate this rank's chunks in time and gather every rank's waveforms on dst."""
        local = torch.cat(self.chunks, dim=-1) if self.chunks else torch.empty(0)
        return gather_rows(local, self.n_total, dst=dst)
