"""Per-CTA event timeline of one tcgen05 GEMM launch (debug hook dph_gemm_debug_timeline): where does the time go?
usage: python tests/gpu_checks/gemm_timeline.py M N K [act] [cta_group] [block_n]"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from dphubert_b200 import _lib, gemm as G  # noqa: E402

dev = torch.device("cuda:0")


def main():
    M, N, K = (int(x) for x in sys.argv[1:4])
    act = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    cg = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    bn = int(sys.argv[6]) if len(sys.argv) > 6 else 0
    a = torch.randn(M, K, device=dev).bfloat16()
    b = torch.randn(N, K, device=dev).bfloat16()
    c = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    bias = torch.randn(N, device=dev) if act else None
    L = _lib.lib()
    buf = torch.zeros(148 * 64, device=dev, dtype=torch.int64)
    run = lambda: G.gemm(G.mat(a), G.mat(b), c, M, N, K, bias=bias, act=act, cta_group=cg, block_n=bn)  # noqa: E731
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    L.dph_gemm_debug_timeline.argtypes = [ctypes.c_void_p]
    assert L.dph_gemm_debug_timeline(ctypes.c_void_p(buf.data_ptr())) == 0
    run()
    torch.cuda.synchronize()
    buf.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run()
    e1.record()
    torch.cuda.synchronize()
    assert L.dph_gemm_debug_timeline(ctypes.c_void_p(0)) == 0
    t = buf.view(148, 64).cpu()
    used = t[:, 0] > 0
    t = t[used]
    t0 = int(t[:, 0].min())
    rel = (t - t0).float() / 1e3  # us
    rel[t == 0] = float("nan")
    rel[:, 48:] = float("nan")
    names = {0: "start", 1: "setup done", 2: "first TMA", 3: "last TMA issued", 4: "role loops done", 5: "end"}
    for i in range(6):
        names[8 + 2 * i] = "tile%d first operands landed" % i
        names[9 + 2 * i] = "tile%d last MMA issued" % i
        names[32 + 2 * i] = "tile%d accumulator ready" % i
        names[33 + 2 * i] = "tile%d epilogue done" % i
    print("GEMM M=%d N=%d K=%d act=%d cg=%d bn=%d: %d CTAs, event time %.1f us, span %.1f us" % (
        M, N, K, act, cg, bn, t.shape[0], e0.elapsed_time(e1) * 1e3, float(torch.nan_to_num(rel).max())))
    print("%-32s %8s %8s %8s  (us after the first CTA started; over CTAs that logged the event)" % ("event", "min", "median", "max"))
    for slot in sorted(names):
        col = rel[:, slot]
        col = col[~torch.isnan(col)]
        if col.numel() == 0:
            continue
        print("%-32s %8.2f %8.2f %8.2f  n=%d" % (names[slot], col.min(), col.median(), col.max(), col.numel()))


    raw = t.float()
    for ti in range(3):
        ld, ep = raw[:, 48 + 4 * ti], raw[:, 49 + 4 * ti]
        m = ld > 0
        if m.any():
            print("tile%d warp2: cycles in tcgen05.ld+wait (sum over chunks) median %.0f, in chunk epilogues median %.0f" % (
                ti, ld[m].median(), ep[m].median()))


if __name__ == "__main__":
    main()
