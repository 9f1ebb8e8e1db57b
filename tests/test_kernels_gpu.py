"""Every C-ABI kernel against its torch restatement (tests/emu.py) on random inputs — runs on the B200 box.
Each case calls the real entry point through ctypes on CUDA tensors, the restatement on CPU copies, and compares every
tensor argument afterwards (outputs are written in place)."""
import pytest
import torch

import emu
from dphubert_b200 import gemm as G
from dphubert_b200 import ops

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")
F32, BF16 = torch.float32, torch.bfloat16


def _cpu(x):
    if torch.is_tensor(x):
        return x.detach().cpu().clone()
    if isinstance(x, (tuple, list)):
        return type(x)(_cpu(v) for v in x)
    if isinstance(x, G.View):
        v = G.View(_cpu(x.t), x.rows, x.K, x.rs, x.cs, x.kin, x.kxs, x.sb0, x.sb1, x.offset)
        return v
    return x


def _flatten(x, out):
    if torch.is_tensor(x):
        out.append(x)
    elif isinstance(x, (tuple, list)):
        for v in x:
            _flatten(v, out)
    elif isinstance(x, dict):
        for v in x.values():
            _flatten(v, out)
    elif isinstance(x, G.View):
        out.append(x.t)


def both(name, args, kwargs=None, tol=None, fn=None, ref=None, ignore=()):
    kwargs = kwargs or {}
    cargs, ckw = _cpu(list(args)), {k: _cpu(v) for k, v in kwargs.items()}
    (fn or getattr(ops, name))(*args, **kwargs)
    torch.cuda.synchronize()
    (ref or getattr(emu, name))(*cargs, **ckw)
    a, b = [], []
    _flatten(list(args), a)
    _flatten(kwargs, a)
    _flatten(cargs, b)
    _flatten(ckw, b)
    assert len(a) == len(b)
    for i, (x, y) in enumerate(zip(a, b)):
        if any(x is t for t in ignore):
            continue
        if not x.is_floating_point():
            assert torch.equal(x.cpu(), y), "%s: integer tensor #%d differs" % (name, i)
            continue
        t = tol if tol is not None else (2e-2 if x.dtype == BF16 else 2e-5)
        xf, yf = x.detach().float().cpu(), y.float()
        assert torch.isfinite(xf).all(), "%s: tensor #%d has non-finite values" % (name, i)
        err = (xf - yf).abs().max().item() / max(yf.abs().max().item(), 1e-6)
        assert err < t, "%s: tensor #%d rel err %.3e (tol %.1e)" % (name, i, err, t)


def rnd(*shape, dtype=F32, scale=1.0):
    return (torch.randn(*shape, device=DEV) * scale).to(dtype)


def zer(*shape, dtype=F32):
    return torch.zeros(*shape, device=DEV, dtype=dtype)


@pytest.fixture(autouse=True)
def _seed():
    torch.manual_seed(1234)
    ops.rng_set(0)


# ---------------------------------------------------------------------------------------------- elementwise / norms
@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("D", [512, 768, 1024, 104])
def test_layernorm_fwd_bwd(dt, D):
    rows = 333
    x, w, b = rnd(rows, D, dtype=dt), rnd(D) * 0.2 + 1, rnd(D) * 0.1
    cm = torch.rand(D, device=DEV)
    for act, colmask in ((0, None), (1, cm)):
        y, ypre, mean, rstd = zer(rows, D, dtype=dt), zer(rows, D, dtype=dt), zer(rows), zer(rows)
        both("layernorm_fwd", (x, w, b, y, rows, D, 1e-5), dict(mean=mean, rstd=rstd, y_pre=ypre, colmask=colmask, act=act))
        dy, dy2 = rnd(rows, D, dtype=dt), rnd(rows, D, dtype=dt)
        dx, dw, db, dm = zer(rows, D, dtype=dt), zer(D), zer(D), zer(D)
        both("layernorm_bwd", (dy, x, w, b, mean, rstd, dx, dw, db, rows, D),
             dict(dy2=dy2, colmask=colmask, dmask=dm if colmask is not None else None, act=act),
             tol=3e-2 if dt == BF16 else 2e-4)
    # padded per-batch output layout (conv dgrad staging)
    Bn, Tn, padr = 3, 111, 3
    dzp = zer(Bn, Tn + 2 * padr, D, dtype=dt)
    dy = rnd(Bn * Tn, D, dtype=dt)
    mean, rstd = rnd(Bn * Tn) * 0.1, torch.rand(Bn * Tn, device=DEV) + 0.5
    xx = rnd(Bn * Tn, D, dtype=dt)
    both("layernorm_bwd", (dy, xx, w, b, mean, rstd, dzp, zer(D), zer(D), Bn * Tn, D),
         dict(rows_per_batch=Tn, out_bs=(Tn + 2 * padr) * D, out_row_off=padr), tol=3e-2 if dt == BF16 else 2e-4)


@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("D", [512, 256, 1024])
def test_layernorm_conv_block_fusions(dt, D):
    """the Large extractor's conv blocks (components.py:54-61, 108-114): y = mask * GELU(LN(x)) and its backward into
    the zero-padded per-utterance row layout — the fused vectorised kernels (D <= 512) and the generic ones (1024)"""
    Bn, Tn, padr = 3, 1111, 3
    rows = Bn * Tn
    x, w, b = rnd(rows, D, dtype=dt), rnd(D) * 0.2 + 1, rnd(D) * 0.1
    cm = torch.rand(D, device=DEV)
    y, mean, rstd = zer(rows, D, dtype=dt), zer(rows), zer(rows)
    both("layernorm_fwd", (x, w, b, y, rows, D, 1e-5), dict(mean=mean, rstd=rstd, colmask=cm, act=1),
         tol=2e-2 if dt == BF16 else 2e-5)
    dy = rnd(rows, D, dtype=dt)
    dzp = zer(Bn, Tn + 2 * padr, D, dtype=dt)
    both("layernorm_bwd", (dy, x, w, b, mean, rstd, dzp, zer(D), zer(D), rows, D),
         dict(colmask=cm, dmask=zer(D), act=1, rows_per_batch=Tn, out_bs=(Tn + 2 * padr) * D, out_row_off=padr),
         tol=3e-2 if dt == BF16 else 2e-4)
    assert float(dzp[:, :padr].abs().max()) == 0.0 and float(dzp[:, -padr:].abs().max()) == 0.0
    # no remap, no mask (conv layer 0 of the Large extractor without a gate)
    dx = zer(rows, D, dtype=dt)
    both("layernorm_bwd", (dy, x, w, b, mean, rstd, dx, zer(D), zer(D), rows, D), dict(act=1),
         tol=3e-2 if dt == BF16 else 2e-4)


@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("D,Dv", [(448, 397), (512, 500), (64, 50)])
def test_layernorm_zero_padded_width(dt, D, Dv):
    """d_valid < D: only the real columns of a zero-padded row are normalised; padding comes back as exact zeros."""
    rows = 257
    x, w, b = rnd(rows, D, dtype=dt), rnd(D) * 0.2 + 1, rnd(D) * 0.1
    for t in (x, w, b):
        t[..., Dv:] = 0
    for act in (0, 1):
        y, mean, rstd = rnd(rows, D, dtype=dt), zer(rows), zer(rows)
        both("layernorm_fwd", (x, w, b, y, rows, D, 1e-5), dict(mean=mean, rstd=rstd, act=act, d_valid=Dv))
        assert float(y[:, Dv:].abs().max()) == 0.0
        dy = rnd(rows, D, dtype=dt)
        dx, dw, db = rnd(rows, D, dtype=dt), zer(D), zer(D)
        both("layernorm_bwd", (dy, x, w, b, mean, rstd, dx, dw, db, rows, D), dict(act=act, d_valid=Dv),
             tol=3e-2 if dt == BF16 else 2e-4)
        assert float(dx[:, Dv:].abs().max()) == 0.0 and float(dw[Dv:].abs().max()) == 0.0


@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("C", [768, 3072, 100])
def test_epilogue_bwd(dt, C):
    rows = 517
    dy, z = rnd(rows, C, dtype=dt), rnd(rows, C, dtype=dt)
    cm, sc = torch.rand(C, device=DEV), torch.rand(1, device=DEV) + 0.5
    tol = 3e-2 if dt == BF16 else 3e-4
    both("epilogue_bwd", (dy, rows, C), dict(z=z, colmask=cm, scalar=sc, dout=zer(rows, C, dtype=dt), dmask=zer(C),
                                             dbias=zer(C), dscalar=zer(1), act=1), tol=tol)
    both("epilogue_bwd", (dy, rows, C), dict(dbias=zer(C)), tol=tol)
    both("epilogue_bwd", (dy, rows, C), dict(z=z, scalar=sc, dout=zer(rows, C, dtype=dt), dbias=zer(C), dscalar=zer(1)),
         tol=tol)
    Bn, Tn, padr = 2, 100, 3
    both("epilogue_bwd", (dy[: Bn * Tn], Bn * Tn, C),
         dict(z=z[: Bn * Tn], colmask=cm, dout=zer(Bn, Tn + 2 * padr, C, dtype=dt), dmask=zer(C), dbias=zer(C), act=1,
              rows_per_batch=Tn, out_bs=(Tn + 2 * padr) * C, out_row_off=padr), tol=tol)


@pytest.mark.parametrize("dt", [F32, BF16])
def test_small_elementwise(dt):
    a, b = rnd(1000, 96, dtype=dt), rnd(1000, 96, dtype=dt)
    both("axpy", (a, b))
    both("axpy", (a.view(-1)[1:], b.view(-1)[1:], 95999))  # unaligned path
    M, H, Dh = 300, 12, 64
    both("group_mask_bwd", (rnd(M, H * Dh, dtype=dt), rnd(M, H * Dh, dtype=dt), torch.rand(H, device=DEV),
                            zer(M, H * Dh, dtype=dt), zer(H), M, H * Dh, Dh), tol=3e-2 if dt == BF16 else 2e-4)
    x = rnd(3, 50, 40, dtype=dt)
    lengths = torch.tensor([50, 20, 0], device=DEV)
    both("pad_rows_zero", (x, lengths, 3, 50, 40))
    for Bn, Tn, Gn, Cg, Cgp, front, pad in ((2, 37, 4, 48, 64, 8, 16), (3, 21, 2, 12, 16, 4, 8)):   # vectorised / scalar
        xp = zer(front + Bn * (Tn + pad) + 16, Gn * Cgp, dtype=dt)
        both("posconv_pad", (rnd(Bn, Tn, Gn * Cg, dtype=dt), xp, Bn, Tn, Gn, Cg, Cgp, front, pad))
    src = rnd(777)
    both("cast", (src, zer(777, dtype=BF16)))
    both("vec_mul", (rnd(100), rnd(100), rnd(100), 0.5, 1.0))


def test_wave_norm():
    wave = rnd(3, 50000) * 0.3 + 0.05
    lengths = torch.tensor([50000, 31234, 1000], device=DEV)
    both("wave_norm", (wave, zer(3, 50000), lengths, 3, 50000), tol=2e-5)
    both("wave_norm", (wave, zer(3, 50000), None, 3, 50000), tol=2e-5)


# ---------------------------------------------------------------------------------------------- conv layer 0
@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("C,bias", [(512, False), (64, True)])
def test_conv0(dt, C, bias):
    Bn, T0, k, s = 2, 16005, 10, 5
    T1 = (T0 - k) // s + 1
    wave = rnd(Bn, T0) * 0.1
    w = rnd(C, 1, k) * 0.3
    bs = rnd(C) * 0.1 if bias else None
    gamma, beta, mask = rnd(C) * 0.2 + 1, rnd(C) * 0.1, torch.rand(C, device=DEV)
    mean, rstd = zer(Bn * C), zer(Bn * C)
    moments = torch.zeros(Bn * 65, device=DEV, dtype=torch.float64)
    ops.conv0_stats(wave, w, bs, moments, mean, rstd, Bn, C, k, s, T1, 1e-5)
    cm, cr = zer(Bn * C).cpu(), zer(Bn * C).cpu()
    emu.conv0_stats(wave.cpu(), w.cpu(), None if bs is None else bs.cpu(), None, cm, cr, Bn, C, k, s, T1, 1e-5)
    assert (mean.cpu() - cm).abs().max() < 1e-5 * max(cm.abs().max().item(), 1e-3)
    assert ((rstd.cpu() - cr).abs() / cr).max() < 1e-4
    tol = 3e-2 if dt == BF16 else 3e-4
    for norm, act in ((1, 1), (0, 1), (0, 0)):
        y = zer(Bn, T1, C, dtype=dt)
        both("conv0_fwd", (wave, w, bs, gamma, beta, mean, rstd, mask, y, Bn, C, k, s, T1, norm, act), tol=tol)
        dy = rnd(Bn, T1, C, dtype=dt)
        scratch = zer(16 * Bn * C)
        # with GroupNorm the conv-bias gradient is analytically zero (the norm removes the mean): skip it there
        both("conv0_bwd", (wave, w, bs, gamma, beta, mean, rstd, moments if norm else None, mask, dy, scratch, zer(C, 1, k),
                           zer(C) if (bias and not norm) else None, zer(C), zer(C), zer(C), Bn, C, k, s, T1, norm, act),
             tol=5e-2 if dt == BF16 else 2e-3, ignore=(scratch, moments))


# ---------------------------------------------------------------------------------------------- attention pieces
@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("T", [199, 64, 499])
def test_softmax(dt, T):
    Bn, H = 2, 3
    Tp = (T + 7) // 8 * 8
    S = rnd(Bn, H, T, Tp) * 2
    lengths = torch.tensor([T, T - 17], device=DEV)
    tol = 2e-2 if dt == BF16 else 1e-4
    P = zer(Bn, H, T, Tp, dtype=dt)
    both("softmax_fwd", (S, Tp, P, None, Tp, lengths, Bn, H, T), tol=tol)
    both("softmax_fwd", (S, Tp, zer(Bn, H, T, Tp, dtype=dt), zer(Bn, H, T, Tp, dtype=dt), Tp, None, Bn, H, T), tol=tol)
    dP = rnd(Bn, H, T, Tp)
    both("softmax_bwd", (P, dP, Tp, zer(Bn, H, T, Tp, dtype=dt), Tp, Bn, H, T), tol=3e-2 if dt == BF16 else 2e-4)


@pytest.mark.parametrize("dt", [F32, BF16])
def test_softmax_wavlm_bias_and_gate(dt):
    Bn, Htot, T, Dh, nb = 2, 4, 70, 32, 32
    heads = [0, 2, 3]
    H = len(heads)
    Tp = (T + 7) // 8 * 8
    x = rnd(Bn, T, Htot * Dh, dtype=dt)
    Wg, bg, cst = rnd(8, Dh) * 0.2, rnd(8) * 0.1, rnd(1, Htot, 1, 1) * 0.1 + 1
    gate, ga, gb = zer(Bn, Htot, T), zer(Bn, Htot, T), zer(Bn, Htot, T)
    both("wavlm_gate_fwd", (x, Wg, bg, cst, gate, ga, gb, Bn, T, Htot, Dh), tol=2e-2 if dt == BF16 else 1e-4)
    rt = torch.randint(0, nb, (2 * T - 1,), device=DEV, dtype=torch.int32)
    emb = rnd(nb, Htot)
    hidx = torch.tensor(heads, device=DEV, dtype=torch.int32)
    wl = (rt, emb, gate, hidx, Htot)
    S = rnd(Bn, H, T, Tp)
    P = zer(Bn, H, T, Tp, dtype=dt)
    tol = 2e-2 if dt == BF16 else 1e-4
    both("softmax_fwd", (S, Tp, P, None, Tp, None, Bn, H, T), dict(wavlm=wl), tol=tol)
    dgate, demb = zer(Bn, Htot, T), zer(nb, Htot)
    both("softmax_bwd", (P, rnd(Bn, H, T, Tp), Tp, zer(Bn, H, T, Tp, dtype=dt), Tp, Bn, H, T),
         dict(wavlm=wl, dgate=dgate, demb=demb, num_buckets=nb), tol=3e-2 if dt == BF16 else 5e-4)
    dx = rnd(Bn, T, Htot * Dh, dtype=dt)
    both("wavlm_gate_bwd", (x, Wg, cst, ga, gb, rnd(Bn, Htot, T), dx, zer(8, Dh), zer(8), zer(1, Htot, 1, 1), Bn, T, Htot,
                            Dh), tol=3e-2 if dt == BF16 else 5e-4)


# ---------------------------------------------------------------------------------------------- loss / gates / optimizer
@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("cos_type", [0, 1])
def test_distill_loss(dt, cos_type):
    rows, D = 1001, 768
    a, b = rnd(rows, D, dtype=dt), rnd(rows, D, dtype=dt)
    rowstat, acc, out = zer(rows * 3), zer(4), zer(16)
    both("distill_loss_fwd", (a, b, rowstat, acc, out, rows, D, 0.5, 1.0, 1.0, cos_type, True, False, 2 * rows), tol=1e-4)
    both("distill_loss_fwd", (b, a, zer(rows * 3), acc, out, rows, D, 0.5, 1.0, 1.0, cos_type, False, True, 2 * rows),
         tol=1e-4)
    both("distill_loss_bwd", (a, b, rowstat, None, zer(rows, D, dtype=dt), rows, 2 * rows, D, 0.5, 1.0, 1.0, cos_type),
         tol=2e-2 if dt == BF16 else 1e-4)


def test_hardconcrete_and_size():
    n = 4000
    la, u = rnd(n) * 2 + 1, torch.rand(n, device=DEV).clamp(1e-6, 1 - 1e-6)
    beta, l, r = 2 / 3, -0.1, 1.1
    mask = zer(n)
    both("hardconcrete_fwd", (la, u, None, mask, n, beta, l, r, 1e-6, 0), tol=1e-5)
    both("hardconcrete_bwd", (la, u, rnd(n), zer(n), n, beta, l, r), tol=1e-4)
    # device-generated noise: right range, right moments, new draw after the epoch advances
    u1, u2 = zer(n), zer(n)
    ops.hardconcrete_fwd(la, None, u1, mask, n, beta, l, r, 1e-6, 77)
    ops.rng_advance()
    ops.hardconcrete_fwd(la, None, u2, mask, n, beta, l, r, 1e-6, 77)
    assert u1.min() >= 1e-6 and u1.max() <= 1 - 1e-6 and abs(u1.mean().item() - 0.5) < 0.03
    assert (u1 != u2).float().mean() > 0.99
    seg_off = torch.tensor([0, 512, 1024, 1040], device=DEV)
    seg_len = torch.tensor([512, 512, 12, 2960], device=DEV, dtype=torch.int32)
    norms = zer(4)
    both("l0_norms", (la, seg_off, seg_len, 4, 1.6, norms), tol=1e-5)
    both("l0_norms_bwd", (la, zer(n), seg_off, seg_len, 4, 1.6, rnd(4), None, torch.tensor([0.3], device=DEV)), tol=1e-4)
    both("lagrangian", (torch.tensor([8.0e7], device=DEV), torch.tensor([0.3, 0.2], device=DEV), 9.4e7,
                        torch.tensor([0.75], device=DEV), zer(8)), tol=1e-5)


def test_adamw_and_sumsq():
    n = 100003
    p, g, m, v = rnd(n), rnd(n) * 0.1, rnd(n) * 0.01, torch.rand(n, device=DEV) * 1e-3
    gn = torch.zeros(1, device=DEV, dtype=torch.float64)
    both("sumsq", (g[:100000], gn), tol=1e-6)
    lr, step = torch.tensor([-0.01], device=DEV), torch.tensor([3], device=DEV, dtype=torch.int32)
    p16 = zer(n, dtype=BF16)
    both("adamw", (p, g, m, v, lr, 0.9, 0.999, 1e-8, 0.01, step, gn, 1.0, p16, True), tol=1e-5, ignore=(p16,))
    assert (p16.float() - p).abs().max() <= p.abs().max() * 2 ** -8
    both("counter_inc", (step,))


@pytest.mark.parametrize("Cout,Cin,k,perm,D,Gn,K", [(64, 128, 3, [1, 2, 0], 192, 4, 16),        # tiled kernels
                                                     (512, 512, 2, [1, 0], 768, 16, 128),         # the real geometry
                                                     (48, 40, 5, [4, 3, 2, 1, 0], 96, 2, 24)])    # general fallbacks
def test_weight_prep(Cout, Cin, k, perm, D, Gn, K):
    w = rnd(Cout, Cin, k)
    both("conv_weight_prep", (w, zer(Cout, k, Cin, dtype=BF16), zer(k, Cin, Cout, dtype=BF16), Cout, Cin, k, perm))
    both("conv_wgrad_unpermute", (rnd(Cout, k * Cin), zer(Cout, Cin, k), Cout, Cin, k))
    Cg, Cgp = 48, 64
    v, g = rnd(D, Cg, K), rnd(1, 1, K) + 2
    norm = zer(K)
    both("posconv_weight_prep", (v, g, norm, zer(D, Cg, K), zer(Gn, Cg, K, Cgp, dtype=BF16), zer(Gn, Cg, K, Cgp, dtype=BF16),
                                 D, Gn, Cg, Cgp, K))
    scratch = zer(K)
    both("posconv_weight_bwd", (rnd(Gn, Cgp, K, Cgp), v, g, norm, scratch, zer(D, Cg, K), zer(1, 1, K), D, Gn, Cg, Cgp, Cgp,
                                K), tol=2e-4, ignore=(scratch,))


# ---------------------------------------------------------------------------------------------- fused attention
@pytest.mark.parametrize("T", [64, 199, 499])
@pytest.mark.parametrize("masked", [False, True])
def test_fused_attention(T, masked):
    Bn, H = 2, 3
    inner = H * 64
    qkv = rnd(Bn * T, 3 * inner, dtype=BF16)
    lengths = torch.tensor([T, max(T - 37, 1)], device=DEV) if masked else None
    hmask = torch.rand(H, device=DEV) if masked else None
    ctx, ctx_un, lse = zer(Bn * T, inner, dtype=BF16), zer(Bn * T, inner, dtype=BF16), zer(Bn, H, T)
    both("attn_fwd", (qkv, ctx, ctx_un if masked else None, lse, Bn, H, T, 0.125),
         dict(lengths=lengths, head_mask=hmask), tol=2e-2)
    dctx = rnd(Bn * T, inner, dtype=BF16)
    o_un = ctx_un if masked else ctx
    delta, dq32 = zer(Bn, H, T), zer(Bn * T, inner)
    dqkv = zer(Bn * T, 3 * inner, dtype=BF16)
    both("attn_bwd", (qkv, dctx, o_un, lse, delta, dq32, dqkv, zer(H) if masked else None, Bn, H, T, 0.125),
         dict(lengths=lengths, head_mask=hmask), tol=3e-2, ignore=(dq32,))
    assert float(dq32.abs().max()) == 0.0  # the fp32 dQ accumulator is handed back cleared


def test_fused_attention_wavlm():
    Bn, Htot, T, nb = 2, 4, 150, 32
    heads = [0, 2, 3]
    H = len(heads)
    inner = H * 64
    qkv = rnd(Bn * T, 3 * inner, dtype=BF16)
    gate = torch.rand(Bn, Htot, T, device=DEV) + 0.5
    rt = torch.randint(0, nb, (2 * T - 1,), device=DEV, dtype=torch.int32)
    emb = rnd(nb, Htot)
    wl = (rt, emb, gate, torch.tensor(heads, device=DEV, dtype=torch.int32), Htot)
    ctx, lse = zer(Bn * T, inner, dtype=BF16), zer(Bn, H, T)
    both("attn_fwd", (qkv, ctx, None, lse, Bn, H, T, 0.125), dict(wavlm=wl), tol=2e-2)
    dq32 = zer(Bn * T, inner)
    both("attn_bwd", (qkv, rnd(Bn * T, inner, dtype=BF16), ctx, lse, zer(Bn, H, T), dq32, zer(Bn * T, 3 * inner, dtype=BF16),
                      None, Bn, H, T, 0.125), dict(wavlm=wl, dgate=zer(Bn, Htot, T), demb=zer(nb, Htot), num_buckets=nb),
         tol=3e-2, ignore=(dq32,))


def test_fused_attention_dropout_matches_unfused_mask():
    """same counter-based keep mask as the unfused softmax kernel: fused O == (dropout(P) V) of the unfused path"""
    Bn, H, T = 1, 2, 130
    inner = H * 64
    qkv = rnd(Bn * T, 3 * inner, dtype=BF16)
    ctx = zer(Bn * T, inner, dtype=BF16)
    ops.attn_fwd(qkv, ctx, None, None, Bn, H, T, 0.125, drop_p=0.2, seed=99)
    x = qkv.view(Bn, T, 3, H, 64).float()
    k, v, q = x[:, :, 0].permute(0, 2, 1, 3), x[:, :, 1].permute(0, 2, 1, 3), x[:, :, 2].permute(0, 2, 1, 3)
    Tp = (T + 7) // 8 * 8
    S = torch.zeros(Bn, H, T, Tp, device=DEV)
    S[..., :T] = 0.125 * q @ k.transpose(-1, -2)
    P, Pd = zer(Bn, H, T, Tp), zer(Bn, H, T, Tp)
    ops.softmax_fwd(S, Tp, P, Pd, Tp, None, Bn, H, T, 0.2, 99)
    ref = (Pd[..., :T] @ v).permute(0, 2, 1, 3).reshape(Bn * T, inner)
    err = (ctx.float() - ref).abs().max().item() / ref.abs().max().item()
    assert err < 2e-2, err


@pytest.mark.parametrize("Bn,T,Gn,Cg,K,mode", [(2, 499, 16, 48, 128, "fwd"), (2, 499, 16, 48, 128, "dgrad"),
                                                (1, 130, 4, 64, 128, "fwd"), (3, 257, 2, 16, 64, "dgrad")])
def test_posconv_tc(Bn, T, Gn, Cg, K, mode):
    """tap-stationary tensor-core positional conv against the gather + einsum restatement (engine geometry: rows
    front + b (T + pad) + t, pad = K, front = K / 2)"""
    Cgp, D = 64, Gn * Cg
    pad, front = K, K // 2
    rows = front + Bn * (T + pad) + K
    x = rnd(Bn * T, D, dtype=BF16)
    xp = zer(rows, Gn * Cgp, dtype=BF16)
    ops.posconv_pad(x, xp, Bn, T, Gn, Cg, Cgp, front, pad)
    wp = zer(Gn, Cg, K, Cgp, dtype=BF16)
    wp[..., :Cg] = rnd(Gn, Cg, K, Cg, dtype=BF16, scale=0.05)
    res = rnd(Bn * T, D, dtype=BF16)
    out = zer(Bn * T, D, dtype=BF16)
    if mode == "fwd":
        both("posconv_tc", (xp, wp, rnd(D), res, out, zer(Bn * T, D, dtype=BF16), Bn, T, D, Gn, Cg, Cgp, K, T + pad,
                            front - K // 2, 1))
    else:
        both("posconv_tc", (xp, wp, None, res, out, None, Bn, T, D, Gn, Cg, Cgp, K, T + pad, front - (K - 1 - K // 2), 0))


def test_conv_out_lengths():
    """frame arithmetic of the conv stack (components.py:116-119) incl. utterances shorter than a kernel"""
    lengths = torch.tensor([160000, 64000, 400, 9, 0, 25, 399, 123457], device=DEV, dtype=torch.int64)
    ks, ss = [10, 3, 3, 3, 3, 2, 2], [5, 2, 2, 2, 2, 2, 2]
    out = ops.conv_out_lengths(lengths, torch.empty_like(lengths), ks, ss)
    ref = emu.conv_out_lengths(lengths.cpu(), torch.empty(8, dtype=torch.int64), ks, ss)
    assert torch.equal(out.cpu(), ref)
    assert out[0].item() == 499 and out[1].item() == 199


def test_attention_dropout_keep_rate():
    """row-key / pair hash of the attention-probability dropout: keep rate ~ 1 - p, independent across rows and
    between the two keys of a pair"""
    Bn, H, T = 2, 4, 256
    S = zer(Bn, H, T, T)
    P, Pd = zer(Bn, H, T, T), zer(Bn, H, T, T)
    ops.softmax_fwd(S, T, P, Pd, T, None, Bn, H, T, 0.1, 7)
    keep = (Pd > 0).float()
    rate = keep.mean().item()
    assert abs(rate - 0.9) < 5e-3, rate
    even, odd = keep[..., 0::2], keep[..., 1::2]
    assert abs(even.mean().item() - odd.mean().item()) < 5e-3
    corr = ((even - 0.9) * (odd - 0.9)).mean().item() / (0.9 * 0.1)
    assert abs(corr) < 2e-2, corr
    rows = keep.reshape(-1, T)
    c2 = ((rows[0::2] - 0.9) * (rows[1::2] - 0.9)).mean().item() / (0.9 * 0.1)
    assert abs(c2) < 2e-2, c2


def test_conv0_full_size_tensor_core_vs_fp32():
    """BASELINE size (8 x 10 s): the bf16 tensor-core forward (hi/lo split taps, tanh-form GELU) against the fp32
    CUDA-core kernel (exact FMA chain, erf GELU) on the same inputs — every frame of every utterance is written — and
    both against torch.nn.functional conv1d + group_norm + gelu on the CPU for (utterance, channel) slices"""
    Bn, T0, k, s, C = 8, 160000, 10, 5, 512
    T1 = (T0 - k) // s + 1
    wave = rnd(Bn, T0) * 0.1
    w = rnd(C, 1, k) * 0.3
    gamma, beta, mask = rnd(C) * 0.2 + 1, rnd(C) * 0.1, torch.rand(C, device=DEV)
    mean, rstd = zer(Bn * C), zer(Bn * C)
    moments = torch.zeros(Bn * 65, device=DEV, dtype=torch.float64)
    ops.conv0_stats(wave, w, None, moments, mean, rstd, Bn, C, k, s, T1, 1e-5)
    y32 = torch.full((Bn, T1, C), float("nan"), device=DEV)
    y16 = torch.full((Bn, T1, C), float("nan"), device=DEV, dtype=BF16)
    ops.conv0_fwd(wave, w, None, gamma, beta, mean, rstd, mask, y32, Bn, C, k, s, T1, 1, 1)
    ops.conv0_fwd(wave, w, None, gamma, beta, mean, rstd, mask, y16, Bn, C, k, s, T1, 1, 1)
    torch.cuda.synchronize()
    assert torch.isfinite(y16.float()).all()
    err = (y16.float() - y32).abs()
    # bf16 rounding of the output (2^-9 relative) + the tanh-form GELU (< 6e-4 absolute)
    assert (err <= 4e-3 * y32.abs() + 1e-3).all(), err.max().item()
    # GroupNorm property at full size: per (utterance, channel) the pre-activation has mean 0 / variance 1, so the
    # statistics the kernel consumed are the ones of the conv output it produced
    z = torch.nn.functional.conv1d(wave[:1, None, :], w, stride=s)[0, :8]
    zn = (z - mean[:8, None]) * rstd[:8, None]
    var = z.var(1, unbiased=False)
    assert zn.mean(1).abs().max() < 1e-3
    assert (zn.var(1, unbiased=False) - var / (var + 1e-5)).abs().max() < 2e-4  # rstd = 1 / sqrt(var + eps)
    # and both kernels against torch on the CPU (conv1d -> per-channel GroupNorm -> erf GELU -> mask), for slices of
    # (utterance, channel block) spread over the batch and the 512 channels (components.py:94-120)
    F = torch.nn.functional
    for bi, c0 in [(0, 0), (3, 252), (7, 508), (5, 129)]:
        cs = slice(c0, c0 + 4)
        zc = F.conv1d(wave[bi].cpu()[None, None, :], w[cs].cpu(), stride=s)
        ref = F.gelu(F.group_norm(zc, 4, gamma[cs].cpu(), beta[cs].cpu(), 1e-5))[0].t() * mask[cs].cpu()
        e32 = (y32[bi, :, cs].cpu() - ref).abs()
        e16 = (y16[bi, :, cs].float().cpu() - ref).abs()
        assert e32.max() < 1e-4 * max(1.0, float(ref.abs().max())), (bi, c0, e32.max().item())
        assert (e16 <= 4e-3 * ref.abs() + 1e-3).all(), (bi, c0, e16.max().item())


def test_conv0_bwd_full_size_tensor_core_vs_fp32():
    """BASELINE size (8 x 10 s, 512 channels): the tensor-core backward (conv0_bwd_tc.cu: recompute + gradient sums
    as two tcgen05 products, du rounded to bf16) against the fp32 CUDA-core kernel fed with the SAME (bf16-valued)
    upstream gradient; a ragged last tile (49 999 frames) and every utterance boundary of the persistent CTAs"""
    Bn, T0, k, s, C = 8, 160000, 10, 5, 512
    T1 = (T0 - k) // s + 1
    wave = rnd(Bn, T0) * 0.1
    w = rnd(C, 1, k) * 0.3
    gamma, beta, mask = rnd(C) * 0.2 + 1, rnd(C) * 0.1, torch.rand(C, device=DEV)
    mean, rstd = zer(Bn * C), zer(Bn * C)
    moments = torch.zeros(Bn * 65, device=DEV, dtype=torch.float64)
    ops.conv0_stats(wave, w, None, moments, mean, rstd, Bn, C, k, s, T1, 1e-5)
    dy16 = (rnd(Bn, T1, C) * (1.0 + rnd(1, 1, C).abs())).to(BF16)
    outs = {}
    for name, dy in (("fp32", dy16.float()), ("tc", dy16)):
        o = [zer(C, 1, k), zer(C), zer(C), zer(C)]   # dW, dgamma, dbeta, dmask
        ops.conv0_bwd(wave, w, None, gamma, beta, mean, rstd, moments, mask, dy, zer(16 * Bn * C), o[0], None, o[1], o[2],
                      o[3], Bn, C, k, s, T1, 1, 1)
        outs[name] = o
    torch.cuda.synchronize()
    for nm, a, b in zip(("dW", "dgamma", "dbeta", "dmask"), outs["tc"], outs["fp32"]):
        assert torch.isfinite(a).all()
        e_max = (a - b).abs().max().item() / b.abs().max().item()
        e_l2 = (a - b).norm().item() / b.norm().item()
        # sums of ~400 000 terms each rounded to 2^-9: measured 1e-3 .. 4e-3
        assert e_max < 2e-2 and e_l2 < 1e-2, (nm, e_max, e_l2)


def test_fused_attention_full_size_vs_torch():
    """BASELINE size (B = 8, H = 12, T = 499): fused tcgen05 attention forward / backward against plain torch fp32
    matmul + softmax + autograd on the same bf16-rounded inputs (tolerance 3e-2 of the tensor's max: bf16 P / dS tiles,
    bf16 dQ reductions)"""
    Bn, H, T, scale = 8, 12, 499, 0.125
    inner = H * 64
    qkv = rnd(Bn * T, 3 * inner, dtype=BF16)
    ctx, lse = zer(Bn * T, inner, dtype=BF16), zer(Bn, H, T)
    ops.attn_fwd(qkv, ctx, None, lse, Bn, H, T, scale)
    x = qkv.view(Bn, T, 3, H, 64).float().detach().requires_grad_(True)
    k, v, q = x[:, :, 0].permute(0, 2, 1, 3), x[:, :, 1].permute(0, 2, 1, 3), x[:, :, 2].permute(0, 2, 1, 3)
    S = scale * q @ k.transpose(-1, -2)
    ref = (torch.softmax(S, -1) @ v).permute(0, 2, 1, 3).reshape(Bn * T, inner)
    err = (ctx.float() - ref).abs().max().item() / ref.abs().max().item()
    assert err < 2e-2, err
    ref_lse = torch.logsumexp(S, -1)
    assert (lse - ref_lse).abs().max().item() < 2e-2
    dctx = rnd(Bn * T, inner, dtype=BF16)
    ref.backward(dctx.float())
    gref = x.grad.reshape(Bn * T, 3 * inner)  # same [k | v | q] layout as dqkv
    delta, dq32 = zer(Bn, H, T), zer(Bn * T, inner)
    dqkv = zer(Bn * T, 3 * inner, dtype=BF16)
    ops.attn_bwd(qkv, dctx, ctx, lse, delta, dq32, dqkv, None, Bn, H, T, scale)
    torch.cuda.synchronize()
    for name, sl in (("dK", slice(0, inner)), ("dV", slice(inner, 2 * inner)), ("dQ", slice(2 * inner, 3 * inner))):
        a, b = dqkv[:, sl].float(), gref[:, sl]
        e = (a - b).abs().max().item() / b.abs().max().item()
        assert e < 3e-2, (name, e)
    # linearity in the upstream gradient (size-independent property of the backward)
    dqkv2 = zer(Bn * T, 3 * inner, dtype=BF16)
    ops.attn_bwd(qkv, (dctx.float() * 2).to(BF16), ctx, lse, delta, dq32, dqkv2, None, Bn, H, T, scale)
    e = (dqkv2.float() - 2 * dqkv.float()).abs().max().item() / dqkv2.float().abs().max().item()
    assert e < 2e-2, e


# ---------------------------------------------------------------------------------------------------------------
# Dropout sites outside attention (all active in the benchmarked step; every parity fixture runs with p = 0):
# the forward kernel and the backward kernel regenerate the mask from (seed, epoch, element index) independently,
# so what has to hold is: identical masks on both sides, keep rate 1 - p, survivors scaled by 1 / (1 - p).
# ---------------------------------------------------------------------------------------------------------------
def _mask_of(with_drop, without, p):
    """keep mask (0 / 1) recovered from an output computed with and without dropout; checks the scale"""
    w, wo = with_drop.float(), without.float()
    big = wo.abs() > 0.05 * wo.abs().mean()
    ratio = (w / wo)[big]
    keep = ratio.abs() > 0.5
    assert (ratio[~keep].abs() < 1e-6).all()
    assert ((ratio[keep] * (1 - p) - 1).abs() < 2e-2).all(), "survivors must be scaled by 1 / (1 - p)"
    full = (w.abs() > 0).float()
    return full, big


@pytest.mark.parametrize("M,N,K", [(3992, 768, 768), (500, 512, 256)])
def test_gemm_epilogue_dropout_matches_backward_mask(M, N, K):
    """out-proj / FFN-out / feature-projection site: GEMM epilogue dropout (components.py:273,742,845) forward,
    `epilogue_bwd` with the same seed backward"""
    p, seed = 0.1, 1234567
    x, w, bias = rnd(M, K, dtype=BF16), rnd(N, K, dtype=BF16, scale=0.05), rnd(N)
    y0, y1 = zer(M, N, dtype=BF16), zer(M, N, dtype=BF16)
    G.gemm(G.mat(x), G.mat(w), y0, M, N, K, bias=bias)
    G.gemm(G.mat(x), G.mat(w), y1, M, N, K, bias=bias, drop_p=p, seed=seed)
    torch.cuda.synchronize()
    fwd_mask, big = _mask_of(y1, y0, p)
    rate = fwd_mask[big].mean().item()
    assert abs(rate - (1 - p)) < 3e-3, rate
    dy = torch.ones(M, N, device=DEV, dtype=BF16)
    g = zer(M, N, dtype=BF16)
    db = zer(N)
    ops.epilogue_bwd(dy, M, N, dout=g, dbias=db, drop_p=p, seed=seed)
    torch.cuda.synchronize()
    bwd_mask = (g.float() > 0).float()
    assert ((g.float()[bwd_mask > 0] * (1 - p) - 1).abs() < 1e-2).all()
    assert torch.equal(bwd_mask[big], fwd_mask[big]), "forward and backward dropout masks differ"
    # bias gradient = column sums of the masked, scaled upstream gradient (fp32 inside the kernel; `g` is bf16-rounded)
    ref_db = bwd_mask.sum(0) / (1 - p)
    assert ((db - ref_db).abs() / ref_db).max().item() < 1e-3
    # a different seed gives a different, equally dense mask
    y2 = zer(M, N, dtype=BF16)
    G.gemm(G.mat(x), G.mat(w), y2, M, N, K, bias=bias, drop_p=p, seed=seed + 1)
    torch.cuda.synchronize()
    m2, _ = _mask_of(y2, y0, p)
    agree = (m2[big] == fwd_mask[big]).float().mean().item()
    assert abs(agree - ((1 - p) ** 2 + p ** 2)) < 5e-3, agree


def test_ffn_intermediate_dropout_forward_backward_consistent():
    """FFN hidden site (components.py:733-737): GEMM epilogue GELU -> dropout -> mask forward; `epilogue_bwd` computes
    dz = dh * dropout * mask * GELU'(z) backward with the same seed"""
    M, F, D, p, seed = 1000, 3072, 768, 0.1, 424242
    x, w, bias = rnd(M, D, dtype=BF16), rnd(F, D, dtype=BF16, scale=0.05), rnd(F)
    cm = torch.rand(F, device=DEV) * 0.5 + 0.5
    h0, h1, z = zer(M, F, dtype=BF16), zer(M, F, dtype=BF16), zer(M, F, dtype=BF16)
    G.gemm(G.mat(x), G.mat(w), h0, M, F, D, bias=bias, act=1, colmask=cm, c2=z)
    G.gemm(G.mat(x), G.mat(w), h1, M, F, D, bias=bias, act=1, colmask=cm, drop_p=p, seed=seed, c2=z)
    torch.cuda.synchronize()
    fwd_mask, big = _mask_of(h1, h0, p)
    assert abs(fwd_mask[big].mean().item() - (1 - p)) < 3e-3
    dh = rnd(M, F, dtype=BF16)
    dz0, dz1 = zer(M, F, dtype=BF16), zer(M, F, dtype=BF16)
    dm0, dm1, db0, db1 = zer(F), zer(F), zer(F), zer(F)
    ops.epilogue_bwd(dh, M, F, z=z, colmask=cm, dout=dz0, dmask=dm0, dbias=db0, act=1)
    ops.epilogue_bwd(dh, M, F, z=z, colmask=cm, dout=dz1, dmask=dm1, dbias=db1, act=1, drop_p=p, seed=seed)
    torch.cuda.synchronize()
    bwd_mask, big2 = _mask_of(dz1, dz0, p)
    both_big = big & big2
    assert torch.equal(bwd_mask[both_big], fwd_mask[both_big]), "forward and backward dropout masks differ"
    # mask gradient: d mask_j = sum_i dh_ij * dropout_ij * GELU(z_ij)
    gz = torch.nn.functional.gelu(z.float())
    ref = (dh.float() * fwd_mask / (1 - p) * gz).sum(0)
    assert (dm1 - ref).abs().max().item() < 2e-2 * ref.abs().max().item()


def test_layernorm_dropout_forward_backward_consistent():
    """Transformer._preprocess site for pre-norm models (components.py:888-891): LayerNorm + dropout fused; the
    backward re-derives the same mask"""
    rows, D, p, seed = 3992, 768, 0.1, 99887766
    x, w, b = rnd(rows, D, dtype=BF16), rnd(D) * 0.1 + 1.0, rnd(D) * 0.1
    y0, y1 = zer(rows, D, dtype=BF16), zer(rows, D, dtype=BF16)
    mean, rstd = zer(rows), zer(rows)
    ops.layernorm_fwd(x, w, b, y0, rows, D, 1e-5, mean, rstd)
    ops.layernorm_fwd(x, w, b, y1, rows, D, 1e-5, mean, rstd, drop_p=p, seed=seed)
    torch.cuda.synchronize()
    fwd_mask, big = _mask_of(y1, y0, p)
    assert abs(fwd_mask[big].mean().item() - (1 - p)) < 3e-3
    dy = rnd(rows, D, dtype=BF16)
    dy_masked = (dy.float() * fwd_mask / (1 - p)).to(BF16)
    dx0, dx1 = zer(rows, D, dtype=BF16), zer(rows, D, dtype=BF16)
    dw0, dw1, db0, db1 = zer(D), zer(D), zer(D), zer(D)
    ops.layernorm_bwd(dy_masked, x, w, b, mean, rstd, dx0, dw0, db0, rows, D)
    ops.layernorm_bwd(dy, x, w, b, mean, rstd, dx1, dw1, db1, rows, D, drop_p=p, seed=seed)
    torch.cuda.synchronize()
    # elements where y0 ~ 0 are excluded from the mask comparison, so compare through the results instead
    err = (dx1.float() - dx0.float()).abs().max().item() / dx0.float().abs().max().item()
    assert err < 3e-2, err
    assert (db1 - db0).abs().max().item() < 2e-2 * db0.abs().max().item() + 0.5
