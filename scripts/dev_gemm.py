"""Dev check of the tcgen05 GEMM through the C ABI (binds vv_linear only)."""
import ctypes as C, sys, time, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
lib = C.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "vibevoice_b200", "libvvb200.so"))
lib.vv_last_error.restype = C.c_char_p
vp, ci = C.c_void_p, C.c_int
lib.vv_linear.argtypes = [vp, ci, vp, ci, ci, ci, ci, ci, vp, ci, vp, vp, vp, ci, vp, ci, ci, ci, vp]
lib.vv_linear_plan.argtypes = [ci, ci, ci, ci, C.POINTER(ci), C.POINTER(ci)]

def run(M, N, K, bn=0, mode=6, splits=0, seed=0, time_it=False):
    g = torch.Generator().manual_seed(seed)
    a = (torch.randn(M, K, generator=g) * 0.5).bfloat16().cuda()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16().cuda()
    bnc, spc = ci(bn), ci(splits)
    lib.vv_linear_plan(M, N, K, mode, C.byref(bnc), C.byref(spc))
    if mode == 4:
        out = torch.zeros(spc.value, M, N, device="cuda")
    else:
        out = torch.zeros(M, N, device="cuda")
    st = vp(torch.cuda.current_stream().cuda_stream)
    def call():
        return lib.vv_linear(vp(a.data_ptr()), K, vp(w.data_ptr()), K, M, N, K, mode, vp(out.data_ptr()), N,
                             None, None, None, 0, None, 0, bnc.value, spc.value, st)
    rc = call()
    torch.cuda.synchronize()
    if rc != 0:
        print("ERR", rc, lib.vv_last_error()); return False
    ref = a.float() @ w.float().t()
    got = out.sum(0) if mode == 4 else out
    err = (got - ref).abs().max().item()
    ok = err < 2e-3 * max(1.0, ref.abs().max().item())
    msg = f"M={M} N={N} K={K} bn={bnc.value} splits={spc.value} mode={mode} err={err:.3e} {'OK' if ok else 'FAIL'}"
    if time_it:
        for _ in range(3): call()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(20): call()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        msg += f"  {ms*1e3:.1f} us  {2*M*N*K/ms/1e9:.1f} TFLOP/s  W {N*K*2/ms/1e6:.0f} GB/s"
    print(msg, flush=True)
    return ok

if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    ok = True
    ok &= run(128, 128, 64, bn=128)
    ok &= run(128, 256, 256, bn=256)
    ok &= run(128, 64, 128, bn=64)
    ok &= run(128, 32, 128, bn=32)
    for (M, N, K) in [(2, 64, 1536), (64, 8192, 2048), (128, 9216, 1536), (128, 1536, 4608), (300, 512, 448),
                      (1000, 128, 32), (4096, 32, 128), (130, 1280, 1024), (77, 96, 72)]:
        for bn in (0, 32, 64, 128, 256):
            ok &= run(M, N, K, bn=bn)
    for splits in (0, 3, 6):
        ok &= run(64, 2048, 8192, mode=4, splits=splits)
        ok &= run(128, 1536, 4608, mode=4, splits=splits)
    print("ALL OK" if ok else "SOME FAILED")
    # timing
    run(8192, 8192, 8192, bn=256, time_it=True)
    run(8192, 8192, 8192, bn=128, time_it=True)
    run(57600, 9216, 1536, bn=256, time_it=True)
    run(128, 9216, 1536, bn=64, time_it=True)
    run(128, 9216, 1536, bn=32, time_it=True)
    run(128, 9216, 1536, bn=128, time_it=True)
    run(128, 21504, 1536, bn=128, time_it=True)
    run(128, 1536, 4608, mode=4, bn=64, splits=6, time_it=True)
    run(128, 1536, 4608, mode=4, bn=128, splits=12, time_it=True)
    run(64, 8192, 2048, bn=64, time_it=True)
    run(64, 2048, 8192, mode=4, bn=64, splits=4, time_it=True)
    run(204800, 128, 32, bn=128, time_it=True)
    run(51200, 512, 128, bn=256, time_it=True)
    sys.exit(0 if ok else 1)
