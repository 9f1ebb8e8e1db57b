"""`dph_conv0_bwd_static` — the allocation-free variant of `dph_conv0_bwd` (scratch in a file-scope device array of the
library; no counterpart in the reference) — called through the C-ABI and
checked against the existing entry point on the same inputs, on two calls in a row (the shared scratch keeps the
previous call's contents, so the second call proves it is cleared).  Collected last on purpose (`-x` runs)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
F32, BF16 = torch.float32, torch.bfloat16


def _setup(dev, Bn, C, T0, dt, seed):
    from dphubert_b200 import ops

    g = torch.Generator(device=dev).manual_seed(seed)

    def rnd(*shape, scale=1.0, dtype=F32):
        return (torch.randn(*shape, device=dev, generator=g) * scale).to(dtype)

    k, s = 10, 5
    T1 = (T0 - k) // s + 1
    wave, w = rnd(Bn, T0, scale=0.1), rnd(C, 1, k, scale=0.3)
    gamma, beta = rnd(C, scale=0.2) + 1, rnd(C, scale=0.1)
    mask = torch.rand(C, device=dev, generator=g)
    mean, rstd = torch.zeros(Bn * C, device=dev), torch.zeros(Bn * C, device=dev)
    moments = torch.zeros(Bn * 65, device=dev, dtype=torch.float64)
    ops.conv0_stats(wave, w, None, moments, mean, rstd, Bn, C, k, s, T1, 1e-5)
    dy = rnd(Bn, T1, C, dtype=dt)
    return dict(wave=wave, w=w, gamma=gamma, beta=beta, mask=mask, mean=mean, rstd=rstd, moments=moments, dy=dy, Bn=Bn,
                C=C, k=k, s=s, T1=T1)


def _outs(dev, C, k):
    return [torch.zeros(C, 1, k, device=dev), torch.zeros(C, device=dev), torch.zeros(C, device=dev),
            torch.zeros(C, device=dev)]   # dW, dgamma, dbeta, dmask


def _rel(a, b):
    return (a - b).abs().max().item() / max(b.abs().max().item(), 1e-12)


@pytest.mark.parametrize("dt", [F32, BF16])
def test_static_scratch_variant_matches_conv0_bwd_twice(dt):
    from dphubert_b200 import ops

    dev = torch.device("cuda:0")
    # two different problems back to back: the second (smaller) one runs on a scratch array still holding the first's
    # per-(b, c) sums and must not see them
    for call, (Bn, C, T0, seed) in enumerate([(3, 512, 16005, 11), (2, 64, 8005, 12)]):
        a = _setup(dev, Bn, C, T0, dt, seed)
        ref, got = _outs(dev, C, a["k"]), _outs(dev, C, a["k"])
        scratch = torch.zeros(16 * Bn * C, device=dev)
        ops.conv0_bwd(a["wave"], a["w"], None, a["gamma"], a["beta"], a["mean"], a["rstd"], a["moments"], a["mask"], a["dy"],
                      scratch, ref[0], None, ref[1], ref[2], ref[3], Bn, C, a["k"], a["s"], a["T1"], 1, 1)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        ops.conv0_bwd_static(a["wave"], a["w"], None, a["gamma"], a["beta"], a["mean"], a["rstd"], a["moments"], a["mask"],
                             a["dy"], got[0], None, got[1], got[2], got[3], Bn, C, a["k"], a["s"], a["T1"], 1, 1)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        # same kernels, same launch order; the per-(b, c) sums are accumulated with atomics, so equal up to their
        # ordering — the tolerance the kernel tests use for this op against its restatement is 2e-3 (fp32) / 5e-2 (bf16)
        tol = 2e-3 if dt == F32 else 5e-2
        for name, x, y in zip(("dW", "dgamma", "dbeta", "dmask"), got, ref):
            assert torch.isfinite(x).all() and _rel(x, y) < tol, (call, name, _rel(x, y))
        assert float(ref[0].abs().max()) > 0


def test_static_scratch_variant_rejects_oversize():
    """B * C beyond the stated maximum is refused with a clear message, before any launch."""
    from dphubert_b200 import _lib

    lib = _lib.lib()
    rc = lib.dph_conv0_bwd_static(None, 0, None, None, None, None, None, None, None, None, None, None, None, None, None,
                                  None, 64, 512, 10, 5, 100, 1, 1, 1, None)
    assert rc < 0
    lib.dph_last_error.restype = ctypes.c_char_p
    msg = lib.dph_last_error().decode()
    assert "exceeds the static scratch limit" in msg and "16384" in msg, msg
