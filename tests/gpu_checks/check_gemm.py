"""Stand-alone GPU check of the tcgen05 GEMM against torch.matmul (run on the B200 box).

    python tests/gpu_checks/check_gemm.py [case-filter]

Each case prints max-abs / relative error; the process exits non-zero if any case fails.
"""
import sys
import time

import torch

sys.path.insert(0, ".")
from dphubert_b200 import gemm as G  # noqa: E402

dev = "cuda"
torch.manual_seed(0)
results = []


def rel_err(out, ref):
    out = out.float()
    ref = ref.float()
    return ((out - ref).abs().max() / (ref.abs().max() + 1e-12)).item()


def report(name, err, tol):
    ok = err == err and err < tol
    results.append((name, ok))
    print("%-58s err=%.3e %s" % (name, err, "OK" if ok else "FAIL"), flush=True)


def rnd(*shape, dtype=torch.bfloat16, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(dtype)


def case_nt(M, N, K, block_n=0, dtype=torch.bfloat16):
    a = rnd(M, K, dtype=dtype)
    b = rnd(N, K, dtype=dtype)
    c = torch.empty(M, N, device=dev, dtype=dtype)
    G.gemm(G.mat(a), G.mat(b), c, M, N, K, block_n=block_n)
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    report("NT %s M=%d N=%d K=%d bn=%d" % (str(dtype)[6:], M, N, K, block_n), rel_err(c, ref), 2e-2 if dtype == torch.bfloat16 else 1e-5)


def case_majors(M, N, K, a_mn, b_mn, block_n=0):
    # a_mn: A stored as (K, M) row-major; b_mn: B stored as (K, N) row-major
    a = rnd(K, M) if a_mn else rnd(M, K)
    b = rnd(K, N) if b_mn else rnd(N, K)
    c = torch.empty(M, N, device=dev, dtype=torch.float32)
    G.gemm(G.mat_t(a) if a_mn else G.mat(a), G.mat_t(b) if b_mn else G.mat(b), c, M, N, K, block_n=block_n)
    torch.cuda.synchronize()
    af = a.float().t() if a_mn else a.float()
    bf = b.float() if b_mn else b.float().t()
    report("majors a_mn=%d b_mn=%d M=%d N=%d K=%d bn=%d" % (a_mn, b_mn, M, N, K, block_n), rel_err(c, af @ bf), 1e-2)


def case_epilogue():
    M, N, K = 500, 320, 192
    a, b = rnd(M, K), rnd(N, K)
    bias = torch.randn(N, device=dev)
    cm = torch.rand(N, device=dev)
    sm = torch.tensor([0.7], device=dev)
    res = rnd(M, N)
    c = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    c2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    G.gemm(G.mat(a), G.mat(b), c, M, N, K, alpha=0.5, bias=bias, act=1, colmask=cm, scalar_mask=sm, residual=res, c2=c2)
    torch.cuda.synchronize()
    z = 0.5 * (a.float() @ b.float().t()) + bias
    ref = torch.nn.functional.gelu(z) * cm * 0.7 + res.float()
    report("epilogue bias+gelu+masks+residual", rel_err(c, ref), 2e-2)
    report("epilogue c2 (pre-activation)", rel_err(c2, z), 2e-2)
    # fp32 accumulate + atomic split-K
    M, N, K = 304, 200, 4096
    a, b = rnd(K, M), rnd(K, N)
    c = torch.ones(M, N, device=dev)
    G.gemm(G.mat_t(a), G.mat_t(b), c, M, N, K, atomic=True, ksplit=8)
    torch.cuda.synchronize()
    report("wgrad-style MN/MN split-K=8 atomic", rel_err(c, 1.0 + a.float().t() @ b.float()), 1e-2)
    c = torch.ones(M, N, device=dev)
    G.gemm(G.mat_t(a), G.mat_t(b), c, M, N, K, accumulate=True)
    torch.cuda.synchronize()
    report("wgrad-style MN/MN accumulate (beta=1)", rel_err(c, 1.0 + a.float().t() @ b.float()), 1e-2)
    # dropout keeps ~ (1-p) of entries and rescales
    M, N, K = 512, 512, 64
    a, b = rnd(M, K), rnd(N, K)
    c = torch.empty(M, N, device=dev)
    G.gemm(G.mat(a), G.mat(b), c, M, N, K, drop_p=0.25, seed=123)
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    kept = c != 0
    frac = kept.float().mean().item()
    e = rel_err(c[kept], ref[kept] / 0.75)
    report("dropout p=.25 keep-frac=%.3f" % frac, e + abs(frac - 0.75), 3e-2)


def case_batched():
    # attention-like: Q,K,V inside a fused (B, T, 3, H, 64) buffer
    B, T, H, Dh = 3, 199, 4, 64
    qkv = rnd(B, T, 3, H, Dh)
    s_t, s_b = 3 * H * Dh, T * 3 * H * Dh
    scores = torch.empty(B, H, T, T, device=dev, dtype=torch.bfloat16)
    q = G.View(qkv, T, Dh, s_t, sb0=Dh, sb1=s_b, offset=0)
    k = G.View(qkv, T, Dh, s_t, sb0=Dh, sb1=s_b, offset=H * Dh)
    G.gemm(q, k, scores, T, T, Dh, nb0=H, nb1=B, c_sb0=T * T, c_sb1=H * T * T, alpha=0.125)
    torch.cuda.synchronize()
    qf, kf, vf = [qkv[:, :, i].float().permute(0, 2, 1, 3) for i in range(3)]
    ref = 0.125 * qf @ kf.transpose(-1, -2)
    report("batched QK^T (strided views of fused qkv)", rel_err(scores, ref), 2e-2)
    # P @ V : B operand MN-major (V is (T, Dh) with T the reduction index)
    p = torch.softmax(ref, -1).to(torch.bfloat16)
    out = torch.empty(B, T, H, Dh, device=dev, dtype=torch.bfloat16)
    pv = G.View(p, T, T, T, sb0=T * T, sb1=H * T * T)
    v = G.View(qkv, Dh, T, 1, cs=s_t, sb0=Dh, sb1=s_b, offset=2 * H * Dh)
    G.gemm(pv, v, out, T, Dh, T, nb0=H, nb1=B, ldc=H * Dh, c_sb0=Dh, c_sb1=T * H * Dh)
    torch.cuda.synchronize()
    ref_o = (p.float() @ vf).permute(0, 2, 1, 3)
    report("batched P@V (B MN-major, K tail 199)", rel_err(out, ref_o), 2e-2)


def case_conv():
    # strided conv as an overlapping-row view of a channels-last activation
    Bn, Tin, C, k, s, Cout = 2, 801, 64, 3, 2, 96
    Tout = (Tin - k) // s + 1
    x = rnd(Bn, Tin, C)
    w = rnd(Cout, C, k, scale=0.1)
    wp = w.permute(0, 2, 1).contiguous().view(Cout, k * C)  # (Cout, tap, Cin)
    y = torch.empty(Bn, Tout, Cout, device=dev, dtype=torch.bfloat16)
    a = G.View(x, Tout, k * C, s * C, sb0=Tin * C)
    G.gemm(a, G.mat(wp), y, Tout, Cout, k * C, nb0=Bn, c_sb0=Tout * Cout)
    torch.cuda.synchronize()
    ref = torch.nn.functional.conv1d(x.float().transpose(1, 2), w.float(), stride=s).transpose(1, 2)
    report("implicit-GEMM conv k=3 s=2 (overlapped rows)", rel_err(y, ref), 2e-2)
    # grouped positional conv: tap-major reduction, per-group batch
    Bn, T, Gn, Cg, kk = 2, 150, 4, 64, 16
    D = Gn * Cg
    pad = kk // 2
    x = rnd(Bn, T, D)
    xp = torch.zeros(Bn, T + 2 * pad, D, device=dev, dtype=torch.bfloat16)
    xp[:, pad:pad + T] = x
    w = rnd(D, Cg, kk, scale=0.1)  # (out, in/groups, k)
    wp = w.view(Gn, Cg, Cg, kk).permute(0, 1, 3, 2).contiguous()  # (g, co, tap, ci)
    y = torch.empty(Bn, T, D, device=dev, dtype=torch.bfloat16)
    Tp = T + 2 * pad
    a = G.View(xp, T, kk * Cg, D, kin=Cg, kxs=D, sb0=Cg, sb1=Tp * D)
    bview = G.View(wp, Cg, kk * Cg, kk * Cg, kin=Cg, kxs=Cg, sb0=Cg * kk * Cg)
    G.gemm(a, bview, y, T, Cg, kk * Cg, nb0=Gn, nb1=Bn, ldc=D, c_sb0=Cg, c_sb1=T * D)
    torch.cuda.synchronize()
    ref = torch.nn.functional.conv1d(x.float().transpose(1, 2), w.float(), padding=pad, groups=Gn)[..., :T].transpose(1, 2)
    report("grouped pos-conv (tap-major K, group batch)", rel_err(y, ref), 2e-2)


def _time(fn, n=20):
    """Device time per call: n calls captured in a CUDA graph (no host launch gaps), replayed 5 times."""
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        fn()
        st.synchronize()
        with torch.cuda.graph(g, stream=st):
            for _ in range(n):
                fn()
    torch.cuda.synchronize()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (5 * n)


def case_perf():
    for (M, N, K) in [(3992, 768, 768), (3992, 2304, 768), (3992, 3072, 768), (3992, 768, 3072), (8192, 8192, 8192),
                      (127992, 512, 1536)]:
        a, b = rnd(M, K), rnd(N, K)
        c = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        bias = torch.randn(N, device=dev)
        for cg in (1, 2):
            for bn in (0, 128, 192, 256):
                ms = _time(lambda: G.gemm(G.mat(a), G.mat(b), c, M, N, K, block_n=bn, cta_group=cg))
                print("perf M=%d N=%d K=%d cg=%d bn=%3d: %.3f ms  %.1f TFLOP/s" % (M, N, K, cg, bn, ms, 2.0 * M * N * K / ms / 1e9), flush=True)
            ms = _time(lambda: G.gemm(G.mat(a), G.mat(b), c, M, N, K, bias=bias, act=1, cta_group=cg))
            print("perf M=%d N=%d K=%d cg=%d bias+gelu: %.3f ms  %.1f TFLOP/s" % (M, N, K, cg, ms, 2.0 * M * N * K / ms / 1e9), flush=True)
        ms = _time(lambda: torch.matmul(a, b.t()))
        print("      cuBLAS reference           : %.3f ms  %.1f TFLOP/s" % (ms, 2.0 * M * N * K / ms / 1e9), flush=True)
    # weight-gradient shapes: MN-major x MN-major, fp32 accumulate, library-chosen split-K
    for (M, N, K) in [(3072, 768, 3992), (768, 3072, 3992), (2304, 768, 3992), (768, 768, 3992), (512, 1536, 127992)]:
        a, b = rnd(K, M), rnd(K, N)
        c = torch.zeros(M, N, device=dev, dtype=torch.float32)
        for cg in (1, 2):
            ms = _time(lambda: G.gemm(G.mat_t(a), G.mat_t(b), c, M, N, K, atomic=True, ksplit=-1, cta_group=cg))
            print("perf wgrad M=%d N=%d K=%d cg=%d: %.3f ms  %.1f TFLOP/s" % (M, N, K, cg, ms, 2.0 * M * N * K / ms / 1e9), flush=True)
        ms = _time(lambda: torch.matmul(a.t(), b))
        print("      cuBLAS reference           : %.3f ms  %.1f TFLOP/s" % (ms, 2.0 * M * N * K / ms / 1e9), flush=True)


def case_epiperf():
    """Cost of each epilogue feature on the FFN-1 shape (graph-timed)."""
    M, N, K = 3992, 3072, 768
    a, b = rnd(M, K), rnd(N, K)
    c = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    c2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    res = rnd(M, N)
    bias = torch.randn(N, device=dev)
    cm = torch.rand(N, device=dev)
    variants = {
        "plain": {}, "bias": dict(bias=bias), "gelu": dict(act=1), "colmask": dict(colmask=cm), "c2": dict(c2=c2),
        "residual": dict(residual=res), "bias+gelu": dict(bias=bias, act=1),
        "bias+gelu+colmask+c2": dict(bias=bias, act=1, colmask=cm, c2=c2), "fp32 out": None,
    }
    for name, kw in variants.items():
        if kw is None:
            cf = torch.empty(M, N, device=dev, dtype=torch.float32)
            ms = _time(lambda: G.gemm(G.mat(a), G.mat(b), cf, M, N, K))
        else:
            ms = _time(lambda: G.gemm(G.mat(a), G.mat(b), c, M, N, K, **kw))
        print("epiperf %-24s %.1f us  %.1f TFLOP/s" % (name, ms * 1e3, 2.0 * M * N * K / ms / 1e9), flush=True)


def case_plain():
    """Plain FFN-1-shaped GEMM launches only (target of `ncu --set full`)."""
    M, N, K = 3992, 3072, 768
    a, b = rnd(M, K), rnd(N, K)
    c = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(6):
        G.gemm(G.mat(a), G.mat(b), c, M, N, K)
    torch.cuda.synchronize()


def case_one():
    """One launch of the FFN-1 GEMM with its full epilogue (target of `ncu --set full`)."""
    M, N, K = 3992, 3072, 768
    a, b = rnd(M, K), rnd(N, K)
    c = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    bias = torch.randn(N, device=dev)
    for _ in range(3):
        G.gemm(G.mat(a), G.mat(b), c, M, N, K, bias=bias, act=1)
        G.gemm(G.mat(a), G.mat(b), c, M, N, K)
    torch.cuda.synchronize()


def main():
    flt = sys.argv[1] if len(sys.argv) > 1 else ""
    cases = {
        "basic": lambda: [case_nt(128, 128, 64, 128), case_nt(128, 64, 64, 64), case_nt(256, 256, 256, 256),
                          case_nt(128, 192, 128, 192)],
        "sizes": lambda: [case_nt(3992, 768, 768), case_nt(1000, 2304, 768), case_nt(100, 72, 40), case_nt(333, 520, 1000),
                          case_nt(4096, 4096, 1024, 256)],
        "f32": lambda: [case_nt(333, 200, 150, dtype=torch.float32)],
        "majors": lambda: [case_majors(256, 256, 256, am, bm, bn) for am in (0, 1) for bm in (0, 1) for bn in (64, 128, 256)]
        + [case_majors(304, 200, 1000, 1, 1), case_majors(768, 3072, 3992, 1, 1), case_majors(1000, 520, 333 * 8, 0, 1)],
        "epilogue": case_epilogue,
        "batched": case_batched,
        "conv": case_conv,
        "perf": case_perf,
        "epiperf": case_epiperf,
        "one": case_one,
        "plain": case_plain,
    }
    for name, fn in cases.items():
        if flt and flt not in name:
            continue
        try:
            fn()
        except Exception as e:  # noqa: BLE001
            results.append((name, False))
            print("CASE %s raised: %r" % (name, e), flush=True)
    bad = [n for n, ok in results if not ok]
    print("SUMMARY: %d/%d ok; failed: %s" % (len(results) - len(bad), len(results), bad))
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
