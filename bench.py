#!/usr/bin/env python
"""bench.py — audio-seconds/second of VibeVoice's per-frame acoustic path (diffusion head +
acoustic decoder), BASELINE.json configs[1]: VibeVoice-1.5B head+decoder, bf16, batch 64.

A "step" is one frame-step of the hot path over the whole batch, exactly what one iteration of
generate()'s loop does for the acoustic path (modeling_vibevoice_inference.py:644-658):
    latent = sample_speech_tokens(cond[n,H], neg[n,H], cfg=1.3)      # 10 DPM-Solver++ steps, CFG
    chunk  = acoustic_tokenizer.decode(latent, cache, sample_indices, use_cache=True)   # [n,1,3200]
One step therefore produces n / 7.5 audio-seconds.  A 60 s utterance is 450 such steps; frames
are sequential in the real model (the LLM sits between them), so throughput is per frame-step.

    python bench.py --gpus N --steps K --warmup W            # this framework (sm_100a kernels)
    python bench.py --impl reference ...                     # the reference algorithm on host cores

Prints ONE JSON line (rank 0).  See DESIGN.md §Measurement for every field.  Next to the metric, measured on rank 0
and NOT part of `value`: `frame_tail` (semantic encode + connectors of the same frame), `frame_loop` (the whole
frame iteration through FrameLoopB200 against the reference's host-side bookkeeping around the same kernels) and
`prefill` (voice-prompt encode + sample + connector); `--no-frame-tail` skips the three.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

FRAME_RATE = 7.5          # Hz, hop 3200 @ 24 kHz (README.md:24 of the reference)
N_PER_GPU = 64            # BASELINE.json configs[1]: batch 64
DDPM_STEPS = 10           # demo/inference_from_file.py:313
CFG_SCALE = 1.3           # demo/inference_from_file.py:174-177
UTTERANCE_S = 60


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops_sustained"]), "source": "measured"}
    # /opt/skills/guides/B200_PROFILING.md fallback
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1400.0, "source": "fallback"}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ------------------------------------------------------------------------------------------
# GEMM records -> roofline of the dominant kernel class
# ------------------------------------------------------------------------------------------
TRAFFIC_JSON = os.path.join(ROOT, "profiles", "r02_traffic.json")
TRAFFIC_SOURCES = ["vv_chain.cu", "vv_chain.cuh", "vv_common.cuh", "vv_gemm_dev.cuh", "vv_head.cu", "vv_dec.cu"]


def kernel_source_hash():
    """sha256[:16] of the sources the persistent kernels are built from (there is no .git on the GPU box)."""
    import hashlib
    h = hashlib.sha256()
    for f in TRAFFIC_SOURCES:
        with open(os.path.join(ROOT, "vibevoice_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def measured_traffic(key, n):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of kernel `key` from the committed `ncu --set full`
    capture (profiles/r02_traffic.json, written by scripts/make_traffic_json.py), or None when the capture was made on
    other kernel sources or another batch size: a stale constant is worse than null."""
    try:
        with open(TRAFFIC_JSON) as f:
            d = json.load(f)
        e = d[key]
        if d.get("source_sha16") != kernel_source_hash() or int(e.get("batch", -1)) != n or DDPM_STEPS != int(d.get("ddpm_steps", -1)):
            return None
        return int(e["dram_bytes"])
    except Exception:
        return None


def gemm_bytes(r) -> int:
    """Algorithmic bytes of one GEMM launch: weights once + A once + output once."""
    out_b = {0: 2, 1: 2, 2: 1, 3: 4, 4: 4 * max(r.splits, 1), 5: 8 + 2, 6: 4, 7: 2}[r.mode]  # per output element
    return r.N * r.K * 2 + r.M * r.K * 2 + r.M * r.N * out_b


def gemm_flops(r) -> int:
    return 2 * r.M * r.N * r.K


# ------------------------------------------------------------------------------------------
# This framework's arm
# ------------------------------------------------------------------------------------------
def run_b200(args):
    import torch.distributed as dist

    from vibevoice_b200 import _lib, synth
    from vibevoice_b200._lib import check, ptr
    from vibevoice_b200.decoder import AcousticDecoderB200
    from vibevoice_b200.head import DiffusionHeadB200

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; vibevoice_b200 has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.batch
    hcfg, dcfg = synth.HEAD_1P5B, synth.DEC_FULL
    lib = _lib.load()

    # random-init weights of the named architecture (no network for checkpoints)
    head = DiffusionHeadB200(synth.init_head(hcfg, seed=1234), hcfg, device=dev, max_rows=2 * n)
    dec = AcousticDecoderB200(synth.init_decoder(dcfg, seed=1234), dcfg, device=dev, max_rows=n)
    head.set_schedule(DDPM_STEPS, "dpmsolver++")
    cache = dec.new_cache(n)
    idx = torch.arange(n, device=dev)
    st = torch.cuda.current_stream()
    stp = C.c_void_p(st.cuda_stream)

    # synthetic per-frame conditions: a pool of P distinct frames, rotated (rank-dependent seed:
    # every GPU decodes its own 64 utterances — rows shard data-parallel, weights replicated)
    P = 8
    g = torch.Generator().manual_seed(4321 + rank)
    cond_h = torch.randn(P, n, hcfg.hidden, generator=g).bfloat16().pin_memory()
    neg_h = torch.randn(P, n, hcfg.hidden, generator=g).bfloat16().pin_memory()
    noise_h = torch.randn(P, 2 * n, hcfg.latent, generator=g).pin_memory()
    cond_d, neg_d, noise_d = cond_h.to(dev), neg_h.to(dev), noise_h.to(dev)
    lat_d = torch.empty(n, hcfg.latent, device=dev, dtype=torch.bfloat16)
    wave_d = torch.empty(n, 1, dcfg.hop, device=dev, dtype=torch.bfloat16)
    wave_h = torch.empty(n, 1, dcfg.hop, dtype=torch.bfloat16).pin_memory()

    def step_resident(i):
        # the two C-ABI calls generate() makes per frame, inputs already in HBM
        j = i % P
        check(lib.vv_head_sample(head._h, ptr(cond_d[j]), ptr(neg_d[j]), ptr(noise_d[j]), None, CFG_SCALE,
                                 ptr(lat_d), n, _lib.VV_DTYPE_BF16, stp), "vv_head_sample")
        check(lib.vv_dec_step(dec._h, cache._c, ptr(lat_d), ptr(idx), n, 1, 1, 1.0, 0.0, ptr(wave_d),
                              _lib.VV_DTYPE_BF16, stp), "vv_dec_step")

    def step_e2e(i):
        # public API with HOST buffers: H2D of this frame's conditions + noise, D2H of the audio chunk
        j = i % P
        c = cond_h[j].to(dev, non_blocking=True)
        ng = neg_h[j].to(dev, non_blocking=True)
        lat = head.sample_speech_tokens(c, ng, CFG_SCALE, noise=noise_h[j])
        wave = dec.decode(lat.unsqueeze(1), cache=cache, sample_indices=idx, use_cache=True)
        wave_h.copy_(wave, non_blocking=True)
        st.synchronize()   # the streamer consumes the chunk on the host (streamer.py:50-55)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = lib.vv_launch_count()
        e0.record()
        for i in range(steps):
            fn(warmup + i)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = lib.vv_launch_count() - l0
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms, launches

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms, launches = timed(step_resident, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e, _ = timed(step_e2e, args.steps, args.warmup)

    audio_s_per_step = world * n / FRAME_RATE
    value = audio_s_per_step * args.steps / (ms / 1e3)
    e2e_value = audio_s_per_step * args.steps / (ms_e2e / 1e3)

    # ---- live roofline of the dominant kernel (GEMM launches bracketed by CUDA events) ----
    roof = None
    if rank == 0:
        peaks = load_peaks()
        lib.vv_timing_enable(1)
        R = 3
        for i in range(R):
            step_resident(i)
        torch.cuda.synchronize()
        buf = (_lib.VVLaunchRec * 8192)()
        cnt = lib.vv_timing_fetch(buf, 8192)
        lib.vv_timing_enable(0)
        recs = [buf[i] for i in range(cnt)]
        chain_recs = [r for r in recs if r.kind == 1]       # persistent head launches (all solver steps of a frame)
        gemm_recs = [r for r in recs if r.kind == 0]
        stream_recs = [r for r in gemm_recs if r.M <= 256]  # weight-streaming GEMM launches (HBM-bound class)
        dense_recs = [r for r in gemm_recs if r.M > 256]
        sb = sum(gemm_bytes(r) for r in stream_recs)
        sms = sum(r.ms for r in stream_recs)
        db, dms = sum(gemm_flops(r) for r in dense_recs), sum(r.ms for r in dense_recs)
        step_ms = ms / args.steps

        def hbm_entry(kernel, nbytes, t_ms, launches):
            ach = nbytes / (t_ms / 1e3) / 1e9 if t_ms > 0 else 0.0
            return {"kernel": kernel, "bound": "hbm", "achieved": round(ach, 1), "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": round(ach / peaks["hbm_gbs"], 4), "peak_source": peaks["source"],
                    "launches_per_step": launches / R, "algorithmic_mb_per_step": round(nbytes / R / 1e6, 1),
                    "ms_per_step_in_kernel": round(t_ms / R, 4), "share_of_step": round((t_ms / R) / step_ms, 3)}

        stream_entry = hbm_entry("vv_gemm_kernel (tcgen05/TMA), weight-streaming launches (M<=256)", sb, sms,
                                 len(stream_recs))
        if chain_recs:
            # dominant kernel: the persistent solver loop (one launch per frame-step, all S steps inside)
            cb, cms = sum(r.bytes for r in chain_recs), sum(r.ms for r in chain_recs)
            roof = hbm_entry("k_chain (persistent DPM-Solver++ loop of the head: tcgen05/TMA GEMM phases + row phases, "
                             "%d phases / %d GEMM phases per launch)" % (chain_recs[0].N, chain_recs[0].K), cb, cms,
                             len(chain_recs))
            # dram bytes of one k_chain launch from the committed ncu capture, null if it is not of these sources
            roof["traffic"] = measured_traffic("k_chain_head", n)
            roof["streaming_gemm"] = stream_entry
            stage_recs = [r for r in recs if r.kind == 2]    # decoder stage 0 as one persistent launch (run_stage_chain)
            if stage_recs:
                roof["decoder_stage_chain"] = hbm_entry(
                    "k_chain (decoder stage 0, 8 Block1D at C = 2048: mixer row phase + FFN1 + FFN2 split-K per block, "
                    "%d phases per launch)" % stage_recs[0].N, sum(r.bytes for r in stage_recs),
                    sum(r.ms for r in stage_recs), len(stage_recs))
                roof["decoder_stage_chain"]["traffic"] = measured_traffic("k_chain_decoder_stage0", n)
        else:
            roof = stream_entry
            roof["traffic"] = None
        roof["dense_gemm"] = {"launches_per_step": len(dense_recs) / R, "ms_per_step": round(dms / R, 4),
                              "tflops": round(db / (dms / 1e3) / 1e12, 1) if dms > 0 else None,
                              "frac_of_bf16_peak": round(db / (dms / 1e3) / 1e12 / peaks["bf16_tflops"], 4) if dms > 0 else None}
        roof["all_gemm_ms_per_step"] = round((sms + dms) / R, 4)

    # ---- the rest of the frame iteration, reported next to the metric, not in it:
    # semantic-tokenizer encode of the chunk just decoded + both SpeechConnectors (summed embedding)
    frame_tail = None
    frame_loop = None
    if rank == 0 and not args.no_frame_tail:
        from vibevoice_b200.encoder import SemanticEncoderB200, SpeechConnectorB200
        ecfg = synth.ENC_FULL
        enc = SemanticEncoderB200(synth.init_encoder(ecfg, seed=1234), ecfg, device=dev, max_rows=n)
        ecache = enc.new_cache(n)
        ca, cs = synth.ConnCfg(hcfg.latent, hcfg.hidden), synth.ConnCfg(ecfg.dim, hcfg.hidden)
        conn_a = SpeechConnectorB200(synth.init_connector(ca, seed=1234, tag="a"), ca, device=dev, max_rows=n)
        conn_s = SpeechConnectorB200(synth.init_connector(cs, seed=1234, tag="s"), cs, device=dev, max_rows=n)
        feat = torch.empty(n, 1, ecfg.dim, device=dev, dtype=torch.bfloat16)
        emb_a = torch.empty(n, hcfg.hidden, device=dev, dtype=torch.bfloat16)
        emb = torch.empty(n, hcfg.hidden, device=dev, dtype=torch.bfloat16)

        def step_tail(i):
            check(lib.vv_enc_step(enc._h, ecache._c, ptr(wave_d), ptr(idx), n, 1, ptr(feat), _lib.VV_DTYPE_BF16, stp), "vv_enc_step")
            check(lib.vv_conn_forward(conn_a._h, ptr(lat_d), n, _lib.VV_DTYPE_BF16, None, ptr(emb_a), stp), "vv_conn_forward")
            check(lib.vv_conn_forward(conn_s._h, ptr(feat), n, _lib.VV_DTYPE_BF16, ptr(emb_a), ptr(emb), stp), "vv_conn_forward")

        for i in range(args.warmup):
            step_tail(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            step_tail(i)
        e1.record()
        torch.cuda.synchronize()
        frame_tail = {"ms_per_step": round(e0.elapsed_time(e1) / args.steps, 4),
                      "what": "semantic-tokenizer encode of the decoded chunk (streaming, own cache) + acoustic and semantic "
                              "SpeechConnector, batch %d on rank 0; not part of `value`" % n}
        # ---- the whole frame iteration without host syncs against the reference's host-side
        # bookkeeping (nonzero + per-row .item() + list append, modeling_vibevoice_inference.py:588, :653-658) driving
        # the SAME kernels; 58 of 64 rows in diffusion mode, a different subset every frame
        import time as _time
        from vibevoice_b200.frameloop import FrameLoopB200
        K, Wm = args.steps, args.warmup
        DIFF_ID, k_act = 7, max(1, (n * 9) // 10)
        gl = torch.Generator().manual_seed(77)
        toks = torch.full((P, n), 3, dtype=torch.int64)
        for j in range(P):
            toks[j, torch.randperm(n, generator=gl)[:k_act]] = DIFF_ID
        toks = toks.to(dev)
        fin = torch.zeros(n, dtype=torch.bool, device=dev)
        nie = torch.zeros(n, 1, hcfg.hidden, device=dev, dtype=torch.bfloat16)
        loop = FrameLoopB200(head, dec, enc, conn_a, conn_s, batch=n, max_frames=K + Wm + 1, speech_diffusion_id=DIFF_ID,
                             cfg_scale=CFG_SCALE, num_steps=DDPM_STEPS)

        def step_free(i):
            j = i % P
            loop.step(toks[j], fin, cond_d[j], neg_d[j], nie, noise=noise_d[j])

        cache2, ecache2 = dec.new_cache(n), enc.new_cache(n)
        chunks = [[] for _ in range(n)]

        def step_host(i):
            j = i % P
            tok = toks[j]
            didx = torch.arange(n, device=dev)[~fin & (tok == DIFF_ID)]
            if didx.numel() > 0:
                k = didx.numel()
                lat = head.sample_speech_tokens(cond_d[j][didx], neg_d[j][didx], CFG_SCALE, noise=noise_d[j][:2 * k])
                wave = dec.decode(lat.unsqueeze(1), cache=cache2, sample_indices=didx, use_cache=True)
                for ii, sidx in enumerate(didx):
                    q = sidx.item()
                    if not fin[q]:
                        chunks[q].append(wave[ii])
                feat = enc.encode(wave, cache=ecache2, sample_indices=didx, use_cache=True).mean
                nie[didx] = conn_s(feat, addend=conn_a(lat.unsqueeze(1)))

        def wall(fn):
            for i in range(Wm):
                fn(i)
            torch.cuda.synchronize()
            t0 = _time.perf_counter()
            for i in range(K):
                fn(i)
            torch.cuda.synchronize()
            return (_time.perf_counter() - t0) * 1e3 / K

        ms_free, ms_host = wall(step_free), wall(step_host)
        frame_loop = {"ms_per_frame_sync_free": round(ms_free, 4), "ms_per_frame_host_indexed": round(ms_host, 4),
                      "rows": n, "active_rows": k_act,
                      "what": "head + decode + ring append + semantic encode + connectors + masked embedding update for one "
                              "frame, wall clock per frame over %d frames on rank 0: FrameLoopB200.step (row set stays on the "
                              "device, all %d rows run) vs the reference's bookkeeping (nonzero, .item() and finished_tags[idx] "
                              "per row, list append) around the same kernels on the %d active rows; not part of `value`" % (K, n, k_act)}
        del loop, cache2, ecache2, chunks
        del enc, ecache, conn_a, conn_s

    # ---- voice-prompt prefill, one-off per utterance, reported next to the metric:
    # acoustic encoder over whole prompts (any length, time-chunked) + sample ('gaussian') with bias/scale + connector
    prefill = None
    if rank == 0 and not args.no_frame_tail:
        from vibevoice_b200.encoder import AcousticEncoderB200, SpeechConnectorB200
        acfg = synth.ACOUSTIC_ENC_FULL
        pn, psec = 8, 10.0
        L = int(psec * 24000) + 1234                                   # not a whole number of frames
        T = -(-L // acfg.hop)
        aenc = AcousticEncoderB200(synth.init_encoder(acfg, seed=4321), acfg, device=dev, max_rows=256)
        ca = synth.ConnCfg(hcfg.latent, hcfg.hidden)
        conn_p = SpeechConnectorB200(synth.init_connector(ca, seed=1234, tag="a"), ca, device=dev, max_rows=pn * T)
        prompt = (0.1 * torch.randn(pn, L, generator=torch.Generator().manual_seed(5))).to(dev, torch.bfloat16)

        def step_prefill():
            eo = aenc.encode(prompt)
            feats = eo.sample_features("gaussian", -0.0703125, 0.1962890625)
            return conn_p(feats)

        step_prefill()
        torch.cuda.synchronize()
        reps = 3
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(reps):
            emb_p = step_prefill()
        e1.record()
        torch.cuda.synchronize()
        pms = e0.elapsed_time(e1) / reps
        assert emb_p.shape == (pn, T, hcfg.hidden) and bool(torch.isfinite(emb_p.float()).all())
        prefill = {"ms": round(pms, 3), "prompts": pn, "prompt_s": round(L / 24000.0, 3), "frames": pn * T,
                   "prompt_audio_s_per_s": round(pn * L / 24000.0 / (pms / 1e3), 1),
                   "what": "acoustic_tokenizer.encode (prompt of %d samples, time-chunked 32 frames per utterance) + "
                           "sample('gaussian') with bias/scale + acoustic SpeechConnector, rank 0; not part of `value`" % L}
        del aenc, conn_p, prompt

    # ---- what a user of the reference gets today on this GPU: its torch path, eager bf16 (rank 0, not part of `value`)
    gpu_torch = None
    if rank == 0 and not args.no_gpu_torch_baseline:
        tp = TorchGpuPath(n, dev)
        tms = tp.run(steps=5, warmup=2)
        gpu_torch = {"value": round((n / FRAME_RATE) / (tms / 1e3), 1), "unit": "audio-s/s", "ms_per_step": round(tms, 3),
                     "dtype": "bf16", "kind": "port",
                     "what": "the reference's torch path for the same frame-step (batch %d, 10 solver steps with CFG, streaming "
                             "decode through the reference's per-(layer, sample) cache dict), eager on cuda:0, restated in "
                             "bench.py (TorchGpuPath; /root/reference does not travel); 2 warm-up + 5 timed steps, CUDA events" % n}
        del tp
        torch.cuda.empty_cache()

    # ---- BASELINE.json configs 2 (offline), 4 and 5, short runs inside the same process so they carry the clock record
    other = None
    if rank == 0 and world == 1 and not args.no_other_configs:
        import importlib.util
        spec = importlib.util.spec_from_file_location("measure_configs", os.path.join(ROOT, "scripts", "measure_configs.py"))
        mc = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mc)
        del head, dec, cache
        torch.cuda.empty_cache()
        sampler2 = ClockSampler(local)
        sampler2.start()
        other = mc.measure(dev)
        other["clocks"] = sampler2.stop()

    # "no NCCL except a final gather" (SURVEY.md section 8(e)): every rank holds the finished waveforms of its 64
    # utterances, [64, 1, 450 * 3200] bf16 = 184 MB (FrameLoopB200.audio()-sized), and rank 0 collects them once
    final_gather_ms = None
    if world > 1:
        full = torch.zeros(n, 1, int(UTTERANCE_S * FRAME_RATE) * dcfg.hop, device=dev, dtype=torch.bfloat16)
        gathered = [torch.empty_like(full) for _ in range(world)] if rank == 0 else None
        dist.gather(full, gathered, dst=0)                               # warm-up (communicator set-up)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        dist.gather(full, gathered, dst=0)
        g1.record()
        barrier()
        t = torch.tensor([g0.elapsed_time(g1)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        final_gather_ms = {"ms": round(t.item(), 3), "bytes_per_rank": full.numel() * 2,
                           "what": "dist.gather of [%d, 1, %d] bf16 per rank to rank 0 (NCCL), max over ranks; once per "
                                   "%d-frame utterance batch, i.e. %.4f ms per frame-step" %
                                   (n, full.shape[-1], int(UTTERANCE_S * FRAME_RATE), t.item() / (UTTERANCE_S * FRAME_RATE))}
        del full, gathered

    out = None
    if rank == 0:
        out = {
            "metric": "audio-sec/sec (head+decoder, 1.5B, batch 64)", "value": round(value, 1), "unit": "audio-s/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(n, world),
            "e2e": {"value": round(e2e_value, 1), "unit": "audio-s/s", "ms_per_step": round(ms_e2e / args.steps, 4),
                    "h2d_bytes_per_step": int(2 * n * hcfg.hidden * 2 + 2 * n * hcfg.latent * 4),
                    "d2h_bytes_per_step": int(n * dcfg.hop * 2)},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roof,
            "frame_tail": frame_tail,
            "frame_loop": frame_loop,
            "prefill": prefill,
            "gpu_torch_baseline": gpu_torch,
            "other_configs": other,
            "final_gather": final_gather_ms,
        }
    if world > 1:
        dist.destroy_process_group()
    return out


# ------------------------------------------------------------------------------------------
# Reference arm / CPU baseline: the reference's algorithm for this path on the host cores.
# The reference is pure Python/PyTorch and /root/reference does not exist on the GPU box, so the
# timed code is oracle/vv_oracle.py — the restatement pinned against the live reference in
# tests/test_oracle_vs_reference.py (kind = "port").  fp32, all host threads.
# ------------------------------------------------------------------------------------------
class CpuPath:
    def __init__(self, n: int, threads: int):
        from oracle import vv_oracle as O
        from vibevoice_b200 import synth

        torch.set_num_threads(threads)
        self.O, self.n, self.threads = O, n, threads
        self.hcfg, self.dcfg = synth.HEAD_1P5B, synth.DEC_FULL
        self.Wh = synth.init_head(self.hcfg, seed=1234)
        self.Wd = synth.init_decoder(self.dcfg, seed=1234)
        self.state = O.DecoderState(n)
        self.idx = torch.arange(n)
        g = torch.Generator().manual_seed(4321)
        self.cond = torch.randn(n, self.hcfg.hidden, generator=g)
        self.neg = torch.randn(n, self.hcfg.hidden, generator=g)
        self.noise = torch.randn(2 * n, self.hcfg.latent, generator=g)

    @torch.no_grad()
    def step(self):
        O = self.O
        lat = O.sample_speech_tokens(self.Wh, self.cond, self.neg, CFG_SCALE, self.noise, steps=DDPM_STEPS,
                                     t_dtype=torch.float32)
        return O.decode(self.Wd, lat.unsqueeze(1), self.dcfg.depths, self.dcfg.ratios, self.dcfg.vae_dim,
                        self.dcfg.eps, state=self.state, idx=self.idx)

    def run(self, steps: int, warmup: int):
        for _ in range(warmup):
            self.step()
        t0 = time.perf_counter()
        for _ in range(steps):
            self.step()
        dt = time.perf_counter() - t0
        return dt / steps


# ------------------------------------------------------------------------------------------
# gpu_torch_baseline: what a user of the reference gets TODAY on this GPU - the reference's torch path for the same
# frame-step, eager, bf16, on cuda (SURVEY.md line 20 / BASELINE.md section 3: "the bar is the reference torch path on
# the same B200").  /root/reference does not exist on the GPU box, so the loop is restated here op for op with plain
# torch calls (same kernels the reference's nn.Modules launch: F.linear, F.conv1d, F.conv_transpose1d, eager
# elementwise), including the reference's own streaming-cache bookkeeping (a dict keyed by (layer, sample) walked with
# sample_indices.tolist(), modular_vibevoice_tokenizer.py:193-233 - a host sync per cached layer) and its per-frame
# scheduler.set_timesteps on the host (schedule/dpm_solver.py:321-423).  Reported next to `value`, never part of it.
# ------------------------------------------------------------------------------------------
class TorchGpuPath:
    def __init__(self, n, dev, hcfg=None, dcfg=None, dtype=torch.bfloat16, seed=1234):
        import torch.nn.functional as F
        from vibevoice_b200 import schedule, synth
        self.F, self.n, self.dev = F, n, dev
        self.hcfg, self.dcfg = hcfg or synth.HEAD_1P5B, dcfg or synth.DEC_FULL
        bf = dtype
        self.Wh = {k: v.to(dev, bf) for k, v in synth.init_head(self.hcfg, seed=seed).items()}
        self.Wd = {k: v.to(dev, bf) for k, v in synth.init_decoder(self.dcfg, seed=seed).items()}
        self.schedule = schedule
        g = torch.Generator().manual_seed(4321)
        self.cond = torch.randn(n, self.hcfg.hidden, generator=g).to(dev, bf)
        self.neg = torch.randn(n, self.hcfg.hidden, generator=g).to(dev, bf)
        self.cache = {}
        self.idx = torch.arange(n, device=dev)
        half = self.hcfg.freq_dim // 2
        self.freqs = torch.exp(-math.log(10000.0) * torch.arange(0, half, dtype=torch.float32, device=dev) / half)

    # -- VibeVoiceTokenizerStreamingCache.get / .set (modular_vibevoice_tokenizer.py:198-233)
    def _get(self, layer, idx):
        states = []
        for i in idx.tolist():
            st = self.cache.get((layer, i))
            if st is None:
                return None
            states.append(st)
        return torch.stack(states, dim=0)

    def _set(self, layer, idx, states):
        for k, i in enumerate(idx.tolist()):
            self.cache[(layer, i)] = states[k].detach()

    def _sconv(self, x, w, b, groups, name):          # SConv1d._forward_streaming (:327-382)
        ctx = w.shape[-1] - 1
        h = self._get(name, self.idx)
        if h is None:
            h = torch.zeros(x.shape[0], x.shape[1], ctx, device=x.device, dtype=x.dtype)
        xin = torch.cat([h, x], dim=2)
        self._set(name, self.idx, xin[:, :, -ctx:])
        return self.F.conv1d(xin, w, b, groups=groups)

    def _sconvtr(self, x, w, b, stride, name):        # SConvTranspose1d._forward_streaming (:478-549)
        k = w.shape[-1]
        T = x.shape[2]
        h = self._get(name, self.idx)
        if h is None:
            h = torch.zeros(x.shape[0], x.shape[1], 0, device=x.device, dtype=x.dtype)
        xin = torch.cat([h, x], dim=2)
        y = self.F.conv_transpose1d(xin, w, b, stride=stride)
        y = y[..., : y.shape[-1] - (k - stride)]
        self._set(name, self.idx, xin[:, :, -(k - 1):])
        return y[..., -T * stride:]

    @staticmethod
    def _rms(x, w, eps):                              # RMSNorm.forward (modular_vibevoice_diffusion_head.py:31-38)
        xf = x.float()
        o = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).type_as(x)
        return o * w if w is not None else o

    def _head(self, z, t, cond):                      # VibeVoiceDiffusionHead.forward (:254-280)
        F, W = self.F, self.Wh
        x = F.linear(z, W["noisy_images_proj.weight"])
        args = t[:, None].float() * self.freqs[None]
        e = torch.cat([torch.cos(args), torch.sin(args)], dim=-1).to(z.dtype)
        temb = F.linear(F.silu(F.linear(e, W["t_embedder.mlp.0.weight"])), W["t_embedder.mlp.2.weight"])
        c = F.linear(cond, W["cond_proj.weight"]) + temb
        sc = F.silu(c)
        for i in range(self.hcfg.layers):
            p = f"layers.{i}."
            shift, scale, gate = F.linear(sc, W[p + "adaLN_modulation.1.weight"]).chunk(3, dim=-1)
            y = self._rms(x, W[p + "norm.weight"], self.hcfg.eps) * (1 + scale) + shift
            hmid = F.silu(F.linear(y, W[p + "ffn.gate_proj.weight"])) * F.linear(y, W[p + "ffn.up_proj.weight"])
            x = x + gate * F.linear(hmid, W[p + "ffn.down_proj.weight"])
        shift, scale = F.linear(sc, W["final_layer.adaLN_modulation.1.weight"]).chunk(2, dim=-1)
        return F.linear(self._rms(x, None, self.hcfg.eps) * (1 + scale) + shift, W["final_layer.linear.weight"])

    @torch.no_grad()
    def step(self):
        n, dev = self.n, self.dev
        # sample_speech_tokens (modeling_vibevoice_inference.py:712-725); set_timesteps runs on the host every frame
        self.schedule.timesteps_and_sigmas.cache_clear()
        ts, sig = self.schedule.timesteps_and_sigmas(DDPM_STEPS)
        condition = torch.cat([self.cond, self.neg], dim=0)
        speech = torch.randn(2 * n, self.hcfg.latent).to(condition)          # drawn on the CPU, like :716
        m_prev, lam_prev = None, None
        for i in range(DDPM_STEPS):
            half = speech[:n]
            combined = torch.cat([half, half], dim=0)
            tt = torch.full((2 * n,), float(ts[i]), device=dev).to(combined)
            eps = self._head(combined, tt, condition)
            c_eps, u_eps = eps[:n], eps[n:]
            half_eps = u_eps + CFG_SCALE * (c_eps - u_eps)
            eps = torch.cat([half_eps, half_eps], dim=0)
            # scheduler.step, dpmsolver++ midpoint, v-prediction (schedule/dpm_solver.py:528-594, 676-764, 935-1022)
            s0, s1 = float(sig[i]), float(sig[i + 1])
            a0, a1 = 1.0 / math.sqrt(1 + s0 * s0), 1.0 / math.sqrt(1 + s1 * s1)
            sg0, sg1 = s0 * a0, s1 * a1
            lam0 = math.log(a0) - math.log(sg0)
            lam1 = (math.log(a1) - math.log(sg1)) if s1 > 0 else float("inf")
            x0 = (a0 * speech.float() - sg0 * eps.float())
            h = lam1 - lam0
            em = math.expm1(-h) if math.isfinite(h) else -1.0
            first = (i == 0) or (i == DDPM_STEPS - 1)
            if first:
                xn = (sg1 / sg0) * speech.float() - a1 * em * x0
            else:
                r0 = (lam0 - lam_prev) / h
                d1 = (1.0 / r0) * (x0 - m_prev)
                xn = (sg1 / sg0) * speech.float() - a1 * em * x0 - 0.5 * a1 * em * d1
            m_prev, lam_prev = x0, lam0
            speech = xn.to(condition.dtype)
        lat = speech[:n].unsqueeze(1)
        # acoustic_tokenizer.decode(latent, cache, sample_indices, use_cache=True) (:1099-1108; TokenizerDecoder.forward :914-951)
        F, W, d = self.F, self.Wd, self.dcfg
        x = lat.permute(0, 2, 1)
        x = self._sconv(x, W["upsample_layers.0.0.conv.conv.weight"], W["upsample_layers.0.0.conv.conv.bias"], 1, "stem")
        for i in range(len(d.depths)):
            if i > 0:
                p = f"upsample_layers.{i}.0.convtr.convtr."
                x = self._sconvtr(x, W[p + "weight"], W[p + "bias"], d.ratios[i - 1], f"up{i}")
            for j in range(d.depths[i]):
                p = f"stages.{i}.{j}."
                C = x.shape[1]
                xn = self._rms(x.transpose(1, 2), W[p + "norm.weight"], d.eps).transpose(1, 2)
                y = self._sconv(xn, W[p + "mixer.conv.conv.conv.weight"], W[p + "mixer.conv.conv.conv.bias"], C, f"dw{i}.{j}")
                x = x + y * W[p + "gamma"].unsqueeze(-1)
                yn = self._rms(x.transpose(1, 2), W[p + "ffn_norm.weight"], d.eps)
                y = F.linear(F.gelu(F.linear(yn, W[p + "ffn.linear1.weight"], W[p + "ffn.linear1.bias"])),
                             W[p + "ffn.linear2.weight"], W[p + "ffn.linear2.bias"]).permute(0, 2, 1)
                x = x + y * W[p + "ffn_gamma"].unsqueeze(-1)
        return self._sconv(x, W["head.conv.conv.weight"], W["head.conv.conv.bias"], 1, "head")

    def run(self, steps, warmup):
        for _ in range(warmup):
            self.step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = self.step()
        e1.record()
        torch.cuda.synchronize()
        assert out.shape == (self.n, 1, self.dcfg.hop) and bool(torch.isfinite(out.float()).all())
        return e0.elapsed_time(e1) / steps


def cpu_sample_text(n):
    same = "the GPU arm's batch" if n == N_PER_GPU else f"batch {n} instead of {N_PER_GPU}"
    return (f"{n} utterances x 1 frame-step per step (head: 10 DPM-Solver++ steps with CFG, then streaming decode "
            f"with cache), fp32, all host threads — the same frame-step as the GPU arm at {same}; the sample is "
            f"bounded in steps (a 60 s utterance is 450 of them)")


def workload_config(n, world):
    """The workload both arms report (BASELINE.json configs[1])."""
    return {"workload": "VibeVoice-1.5B head+decoder bf16, batch 64 utterances of 60 s, streaming frame-steps",
            "batch_per_gpu": n, "global_batch": world * n, "ddpm_steps": DDPM_STEPS, "cfg_scale": CFG_SCALE,
            "step": "1 frame for every utterance (sample_speech_tokens + decode with cache) = n/7.5 audio-s",
            "utterance_s": UTTERANCE_S, "steps_per_utterance": int(UTTERANCE_S * FRAME_RATE),
            "l2": "weights streamed per step (1.77 GB head: 10 x 170 MB + 66 MB adaLN once per frame; 0.69 GB decoder) exceed the 126 MB L2",
            "parallelism": f"dp{world} (rows sharded, weights replicated, no collective in the loop)"}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return None
    threads = os.cpu_count() or 1
    n = args.ref_batch
    cpu = CpuPath(n, threads)
    sec = cpu.run(args.steps, args.warmup)
    value = (n / FRAME_RATE) / sec
    return {
        "impl": "reference",
        "metric": "audio-sec/sec (head+decoder, 1.5B, batch 64)", "value": round(value, 4), "unit": "audio-s/s",
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(sec * 1e3, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        # same workload as the B200 arm; each timed step is a bounded sample of it (cpu_baseline.sample): the
        # frame-step is row-wise independent, so audio-s/s of `n` rows on all host cores is the CPU rate
        "config": workload_config(args.batch, int(os.environ.get("WORLD_SIZE", "1"))),
        "reference_sample_rows": n,
        "reference_dtype": "fp32 on host cores (the reference's CPU path, BASELINE.json configs[0])",
        "cpu_baseline": {"value": round(value, 4), "unit": "audio-s/s", "cores": threads, "kind": "port",
                         "sample": cpu_sample_text(n)},
        "e2e": {"value": round(value, 4), "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=N_PER_GPU, help="utterances per GPU")
    ap.add_argument("--ref-batch", type=int, default=N_PER_GPU, help="utterances per step of the CPU reference (default: the full batch)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-frame-tail", action="store_true", help="skip the semantic encoder + connector timing (rank 0)")
    ap.add_argument("--no-gpu-torch-baseline", action="store_true", help="skip the reference's eager torch path on the GPU (rank 0)")
    ap.add_argument("--no-other-configs", action="store_true", help="skip BASELINE.json configs 2 (offline), 4 and 5 (N = 1 only)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    rank, world, _ = dist_env()
    if args.impl == "reference":
        out = run_reference(args)
    else:
        out = run_b200(args)
        if out is not None and world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            n = args.ref_batch
            sec = CpuPath(n, threads).run(steps=2, warmup=1)
            out["cpu_baseline"] = {"value": round((n / FRAME_RATE) / sec, 4), "unit": "audio-s/s", "cores": threads,
                                   "kind": "port", "sample": cpu_sample_text(n) + "; 1 warm-up + 2 timed steps"}
            # BASELINE.json configs[0] (the reference's own CPU-runnable case: 1 speaker, 10 s script = 75 frames, fp32):
            # a bounded sample of its frame-steps at n = 1, scaled to the 75 frames of the script
            sec1 = CpuPath(1, threads).run(steps=2, warmup=1)
            out["cpu_baseline"]["config1_n1"] = {
                "value": round((1 / FRAME_RATE) / sec1, 4), "unit": "audio-s/s", "s_per_frame": round(sec1, 3),
                "s_for_10s_script_extrapolated": round(75 * sec1, 1), "cores": threads, "kind": "port",
                "sample": "1 utterance x 1 frame-step per step (head + streaming decode), fp32, 1 warm-up + 2 timed steps; "
                          "the 10 s script is 75 such steps (extrapolated, frames are identical work)"}
    if out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
