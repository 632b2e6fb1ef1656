/* vvb200_probe.h - measurement probes behind profiles/r01_mma_issue.md and profiles/r01_tma_stream.md.  NOT part of the
 * product ABI: libvvb200.so exports them only when built with `make -C vibevoice_b200/csrc PROBE=1`
 * (scripts/probe_commit.py, scripts/probe_tma.py).  No counterpart in the reference. */
#ifndef VVB200_PROBE_H_
#define VVB200_PROBE_H_
#include "vvb200.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Measurement aid: clocks per iteration of (mmas x tcgen05.mma 128 x bn x 16, commits x tcgen05.commit) on one thread per CTA;
 * nacc accumulators used round-robin (0: four unrolled MMAs on one accumulator). */
VV_API int vv_debug_mma_commit(int bn, int mmas, int commits, int iters, long long* out_host, int ctas, int nacc);
/* Measurement aid (scripts/probe_tma.py, profiles/): one launch in which `ctas` CTAs each pull their own
 * slab of `buf` (or all the same slab when shared_slab != 0) through a TMA -> mbarrier ring of `stages`
 * boxes of box_rows x 128 bytes and drop it.  flavour 0: 2-D tensor-map loads, 1: 1-D bulk copies. */
VV_API int vv_debug_tma_stream(const void* buf, long long bytes, int box_rows, int stages, int ctas, int flavour,
                               int shared_slab, void* stream);

#ifdef __cplusplus
}
#endif
#endif
