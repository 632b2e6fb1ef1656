"""Sync-free frame iteration over libvvb200.

One iteration of ``generate()`` (modeling_vibevoice_inference.py:586-690) picks the rows in diffusion mode on the
HOST — ``torch.arange(batch)[~finished_tags & (next_tokens == speech_diffusion_id)]`` (:588) is a ``nonzero`` and
therefore a device sync, followed by ``.numel() > 0``, per-row ``sample_idx.item()`` + ``finished_tags[idx]`` + a
Python list append for the audio (:653-658), and the streamer's per-row host work (streamer.py:42-55).  With the
acoustic path at ~2.6 ms per frame those syncs are what the loop waits for.

``FrameLoopB200.step`` does the same work without the row set ever leaving the device: the WHOLE batch runs through
head -> decode -> semantic encode -> connectors at a fixed launch shape (the head and the decoder's GEMMs are
weight-bound, extra rows are nearly free); inactive rows are pointed at a spare cache slot (``vv_frame_select``), their
audio is not appended (``vv_ring_append``) and their embedding is not written (``vv_embed_scatter``).  Per active
row the results equal the index-compacted path the reference takes: rows are independent in every kernel.

    loop = FrameLoopB200(head, dec, enc, conn_a, conn_s, batch=n, max_frames=450, speech_diffusion_id=ID,
                         scaling_factor=s, bias_factor=b, cfg_scale=1.3, num_steps=10)
    for each frame:
        loop.end_speech(next_tokens == speech_end_id)                        # (:556-560) caches of those rows -> 0
        chunk, active = loop.step(next_tokens, finished_tags, pos_hidden, neg_hidden, next_inputs_embeds)
        stream.put(chunk, active)                                            # optional: PinnedChunkStream
    waves = loop.audio()                                                     # one D2H: per-row waveforms (:692-700)

The noise of the solver is drawn for all ``2*batch`` rows on the device (the reference draws ``2*len(active)`` rows on
the CPU, :716 — a count only the host knows); pass ``noise=`` to control it.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from . import _lib
from ._lib import check, current_stream, ptr


class FrameLoopB200:
    def __init__(self, head, dec, enc, conn_a, conn_s, batch: int, max_frames: int, speech_diffusion_id: int,
                 scaling_factor: float = 1.0, bias_factor: float = 0.0, cfg_scale: float = 1.3,
                 num_steps: int = 10, algorithm_type: str = "dpmsolver++", dtype: torch.dtype = torch.bfloat16):
        self.lib = _lib.load()
        self.head, self.dec, self.enc, self.conn_a, self.conn_s = head, dec, enc, conn_a, conn_s
        self.n, self.cap = int(batch), int(max_frames)
        self.device = dec.device
        self.dtype = dtype
        self.io = _lib.VV_DTYPE_BF16 if dtype == torch.bfloat16 else _lib.VV_DTYPE_F32
        if dtype not in (torch.bfloat16, torch.float32):
            raise ValueError("dtype must be bfloat16 or float32")
        self.token_id = int(speech_diffusion_id)
        self.scale, self.bias = float(scaling_factor), float(bias_factor)
        self.cfg_scale = float(cfg_scale)
        self.hop = dec.cfg.hop
        if enc.cfg.hop != self.hop:
            raise ValueError("encoder and decoder hops differ")
        head.set_schedule(num_steps, algorithm_type)
        n = self.n
        self.trash = n                                       # spare slot n: where inactive rows read and write
        self.acoustic_cache = dec.new_cache(n + 1)
        self.semantic_cache = enc.new_cache(n + 1)
        self.acoustic_cache.bind(dec)
        self.semantic_cache.bind(enc)
        d = self.device
        with torch.cuda.device(d):
            self.slots = torch.empty(n, dtype=torch.int64, device=d)
            self.active = torch.zeros(n, dtype=torch.bool, device=d)
            self.compact = torch.empty(n, dtype=torch.int64, device=d)
            self.count = torch.zeros(1, dtype=torch.int32, device=d)
            self.pos = torch.zeros(n, dtype=torch.int32, device=d)
            self.ring = torch.zeros(n, self.cap, self.hop, dtype=dtype, device=d)
            self._zslots = torch.empty(n, dtype=torch.int64, device=d)
        self.frames = 0

    # ---- cache resets without a host round trip -------------------------------------------------------------
    def end_speech(self, mask: torch.Tensor) -> None:
        """acoustic_cache.set_to_zero / semantic_cache.set_to_zero for the rows where ``mask`` (device bool [n]) is
        set (:556-560 does ``nonzero`` + ``numel() > 0`` first).  Rows outside the mask zero the spare slot."""
        m = mask.to(self.device, torch.bool).contiguous()
        inv = ~m                                               # vv_frame_select(tokens=NULL): active = !finished
        with torch.cuda.device(self.device):
            st = current_stream()                              # the stream of THIS device (the caller's current device may differ)
            check(self.lib.vv_frame_select(None, ptr(inv), self.n, 0, self.trash, ptr(self._zslots), None, None, None, st),
                  "vv_frame_select")
            check(self.lib.vv_cache_zero(self.acoustic_cache._c, ptr(self._zslots), self.n, st), "vv_cache_zero")
            check(self.lib.vv_cache_zero(self.semantic_cache._c, ptr(self._zslots), self.n, st), "vv_cache_zero")

    # ---- one frame -------------------------------------------------------------------------------------------
    @torch.no_grad()
    def step(self, next_tokens: torch.Tensor, finished_tags: Optional[torch.Tensor], pos_hidden: torch.Tensor,
             neg_hidden: torch.Tensor, next_inputs_embeds: Optional[torch.Tensor] = None,
             noise: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """next_tokens int64 [n], finished_tags bool [n] (or None), pos_hidden / neg_hidden [n, H] (the last hidden
        state of every row, :640-641 without the row gather), next_inputs_embeds [n, 1, H] or [n, H] (updated in place
        for the active rows, :687).  Returns (audio_chunk [n, 1, hop], active bool [n]); rows with active == False hold
        unspecified audio.  No host synchronisation."""
        n, d, lib = self.n, self.device, self.lib
        with torch.cuda.device(d):
            st = current_stream()                              # the stream of THIS device (the caller's current device may differ)
        tok = next_tokens.to(d, torch.int64).contiguous()
        fin = finished_tags.to(d, torch.bool).contiguous() if finished_tags is not None else None
        assert tok.shape == (n,)
        with torch.cuda.device(d):
            check(lib.vv_frame_select(ptr(tok), ptr(fin) if fin is not None else None, n, self.token_id, self.trash,
                                      ptr(self.slots), ptr(self.active), ptr(self.compact), ptr(self.count), st),
                  "vv_frame_select")
        if noise is None:
            noise = torch.randn(2 * n, self.head.cfg.latent, device=d, dtype=torch.float32)
        lat = self.head.sample_speech_tokens(pos_hidden.to(self.dtype), neg_hidden.to(self.dtype), self.cfg_scale,
                                             noise=noise)                                  # [n, L]
        lat3 = lat.unsqueeze(1)
        chunk = self.dec.decode(lat3, cache=self.acoustic_cache, sample_indices=self.slots, use_cache=True,
                                scale=self.scale, bias=self.bias)                          # [n, 1, hop]
        chunk = chunk.contiguous()
        with torch.cuda.device(d):
            check(lib.vv_ring_append(ptr(chunk), ptr(self.active), n, self.hop, self.io, ptr(self.ring), ptr(self.pos),
                                     self.cap, st), "vv_ring_append")
        feat = self.enc.encode(chunk, cache=self.semantic_cache, sample_indices=self.slots, use_cache=True).mean
        emb_a = self.conn_a(lat3)
        emb = self.conn_s(feat, addend=emb_a)                                              # [n, 1, H] bf16
        if next_inputs_embeds is not None:
            dst = next_inputs_embeds
            H = emb.shape[-1]
            if not (dst.is_cuda and dst.is_contiguous() and dst.numel() == n * H
                    and dst.dtype in (torch.bfloat16, torch.float32)):
                raise ValueError("next_inputs_embeds must be a contiguous CUDA bf16/fp32 tensor of n*H elements")
            e = emb.to(torch.bfloat16).contiguous()
            with torch.cuda.device(d):
                check(lib.vv_embed_scatter(ptr(e), ptr(self.active), n, H, ptr(dst),
                                           _lib.VV_DTYPE_BF16 if dst.dtype == torch.bfloat16 else _lib.VV_DTYPE_F32, st),
                      "vv_embed_scatter")
        self.frames += 1
        self.last_latent, self.last_embed = lat3, emb
        return chunk, self.active

    # ---- results ---------------------------------------------------------------------------------------------
    def audio(self) -> List[Optional[torch.Tensor]]:
        """Per-row waveform [1, frames_i * hop] (None for a row that never spoke), like the concatenation of
        ``audio_chunks[i]`` (:692-700).  The one synchronising call of the loop."""
        pos = self.pos.cpu().tolist()
        if max(pos) > self.cap:
            raise _lib.VVError(f"audio ring overflow: a row produced {max(pos)} frames, capacity {self.cap}")
        out = []
        for i, p in enumerate(pos):
            out.append(self.ring[i, :p].reshape(1, p * self.hop) if p > 0 else None)
        return out

    def reset(self) -> None:
        self.pos.zero_()
        self.acoustic_cache.clear()
        self.semantic_cache.clear()
        self.frames = 0


class PinnedChunkStream:
    """Device -> host hand-off of the per-frame audio for a streamer (streamer.py:42-55 does ``.item()`` and a
    ``.cpu()`` per row per frame): ``put`` enqueues ONE async copy of the whole chunk and its active mask into one of
    two pinned buffers and records an event; ``get`` (consumer side) waits on that event only — the producer stream
    is never synchronised.  Double-buffered: ``get`` frame k before ``put`` of frame k+2."""

    def __init__(self, n: int, hop: int, dtype: torch.dtype = torch.bfloat16, depth: int = 2):
        self.bufs = [torch.empty(n, hop, dtype=dtype).pin_memory() for _ in range(depth)]
        self.acts = [torch.empty(n, dtype=torch.bool).pin_memory() for _ in range(depth)]
        self.events = [torch.cuda.Event() for _ in range(depth)]
        self.depth, self.k_put, self.k_get = depth, 0, 0

    def put(self, chunk: torch.Tensor, active: torch.Tensor) -> None:
        if self.k_put - self.k_get >= self.depth:
            raise RuntimeError("PinnedChunkStream: consumer is more than `depth` frames behind")
        s = self.k_put % self.depth
        self.bufs[s].copy_(chunk.reshape(self.bufs[s].shape), non_blocking=True)
        self.acts[s].copy_(active, non_blocking=True)
        self.events[s].record()
        self.k_put += 1

    def get(self):
        """-> (chunk [n, hop] pinned host tensor, active bool [n]) of the oldest frame not yet consumed."""
        if self.k_get >= self.k_put:
            raise RuntimeError("PinnedChunkStream: nothing to get")
        s = self.k_get % self.depth
        self.events[s].synchronize()
        self.k_get += 1
        return self.bufs[s], self.acts[s]
