"""CPU emulation of the C-ABI kernels — TEST INFRASTRUCTURE ONLY (never imported by the product).

`install()` monkey-patches `dphubert_b200.ops` / `dphubert_b200.gemm.gemm` with torch-CPU restatements of what each
`dph_*` entry point is specified to compute, so that the *host-side orchestration* (operand views, strides, buffer
layouts, the order of launches in the forward / backward / optimizer, the flat arenas) can be checked against the
oracle without a GPU.  It also enforces the alignment / stride rules of the TMA path on every bf16 GEMM operand.
The CUDA kernels themselves are verified separately on a B200 (`-m gpu` tests)."""
import math

import torch

from dphubert_b200 import engine, gemm as G, ops, trainer

_SQRT1_2 = 0.7071067811865476


def _flat(t):
    """1-D view of the storage behind `t`, starting at t's first element"""
    n = t.untyped_storage().nbytes() // t.element_size() - t.storage_offset()
    return torch.as_strided(t, (n,), (1,), t.storage_offset())


def _gelu(x):
    return 0.5 * x * (1.0 + torch.erf(x * _SQRT1_2))


def _gelu_grad(x):
    return 0.5 * (1.0 + torch.erf(x * _SQRT1_2)) + x * torch.exp(-0.5 * x * x) * 0.3989422804014327


def _check_bf16_operand(v, nb0, nb1):
    o = G._operand_bf16(v, nb0, nb1)  # raises if neither K-major nor MN-major
    assert (v.offset * 2) % 16 == 0 and (v.t.storage_offset() * 2) % 16 == 0, "operand not 16-byte aligned"
    nd = 5
    for i in range(4):
        if o.dims[i + 1] > 1:
            assert o.strides[i] % 8 == 0 and o.strides[i] > 0, "TMA stride %d = %d not a positive multiple of 8" % (
                i, o.strides[i])
    if not o.major_mn and o.dims[2] > 1:
        assert o.dims[0] % 64 == 0, "split reduction view needs dims[0] % 64 == 0 (got %d)" % o.dims[0]
    assert v.rs >= 0 and v.cs >= 0 and v.kxs >= 0 and v.sb0 >= 0 and v.sb1 >= 0


def _gather(v, nb0, nb1):
    flat = _flat(v.t)
    rows = torch.arange(v.rows).view(1, 1, -1, 1)
    k = torch.arange(v.K).view(1, 1, 1, -1)
    b0 = torch.arange(nb0).view(1, -1, 1, 1)
    b1 = torch.arange(nb1).view(-1, 1, 1, 1)
    idx = v.offset + b0 * v.sb0 + b1 * v.sb1 + rows * v.rs + (k // v.kin) * v.kxs + (k % v.kin) * v.cs
    assert idx.min() >= 0 and idx.max() < flat.numel(), "GEMM operand view reaches outside its buffer (%d..%d of %d)" % (
        idx.min(), idx.max(), flat.numel())
    return flat[idx].float()  # (nb1, nb0, rows, K)


def gemm(a, b, c, M, N, K, *, ldc=None, nb0=1, nb1=1, c_sb0=0, c_sb1=0, c_offset=0, alpha=1.0, bias=None, act=0,
         colmask=None, scalar_mask=None, residual=None, ldr=None, r_sb0=0, r_sb1=0, r_offset=0, c2=None,
         accumulate=False, atomic=False, ksplit=1, drop_p=0.0, seed=0, block_n=0, vec_sb0=0, scalar_sb0=0,
         cta_group=0, act_bwd=False, z=None, ldz=None, z_sb0=0, z_sb1=0, z_offset=0, dbias=None, dmask=None,
         b_taps=0):
    assert drop_p == 0.0, "the emulator runs the dropout-free parity configuration"
    if b_taps:
        # taps folded into N: column n = tap * rows + channel reads the view shifted by tap * sb1
        assert a.t.dtype == torch.bfloat16 and nb1 == 1 and b.rows * b_taps == N and b.rs == 1 and b.rows % 64 == 0
        _check_bf16_operand(b, nb0, b_taps)
        for tap in range(b_taps):
            bt = G.View(b.t, b.rows, b.K, b.rs, b.cs, b.kin, b.kxs, b.sb0, 0, b.offset + tap * b.sb1)
            gemm(a, bt, c, M, b.rows, K, ldc=ldc, nb0=nb0, nb1=1, c_sb0=c_sb0, c_sb1=0,
                 c_offset=c_offset + tap * b.rows, alpha=alpha, accumulate=accumulate, atomic=atomic, ksplit=ksplit)
        return c
    ldc = N if ldc is None else ldc
    ldr = ldc if ldr is None else ldr
    assert a.rows == M and b.rows == N and a.K == K and b.K == K, "view extents disagree with the problem size"
    assert a.t.dtype == b.t.dtype
    if a.t.dtype == torch.bfloat16:
        _check_bf16_operand(a, nb0, nb1)
        _check_bf16_operand(b, nb0, nb1)
        if atomic:
            assert c.dtype == torch.float32
        assert ksplit == 1 or atomic
    else:
        assert c.dtype == torch.float32
    A = _gather(a, nb0, nb1)
    Bm = _gather(b, nb0, nb1)
    v = alpha * torch.matmul(A, Bm.transpose(-1, -2))  # (nb1, nb0, M, N)
    b0 = torch.arange(nb0).view(1, -1, 1, 1)
    b1 = torch.arange(nb1).view(-1, 1, 1, 1)
    m = torch.arange(M).view(1, 1, -1, 1)
    n = torch.arange(N).view(1, 1, 1, -1)
    if bias is not None:
        v = v + _flat(bias)[b0 * vec_sb0 + n]
    cidx = (c_offset + b0 * c_sb0 + b1 * c_sb1 + m * ldc + n).expand(nb1, nb0, M, N)
    if act_bwd:
        # fused backward of o = colmask * act(z): c = acc * colmask * act'(z); dbias += colsum(c); dmask += colsum(acc * act(z))
        assert a.t.dtype == torch.bfloat16 and c.dtype == torch.bfloat16 and z.dtype == torch.bfloat16
        assert G.act_bwd_ok(c, z, N, ldc, ldc if ldz is None else ldz, (c_sb0, c_sb1, z_sb0, z_sb1), (c_offset, z_offset))
        assert bias is None and residual is None and c2 is None and scalar_mask is None and alpha == 1.0
        assert not (atomic or accumulate)
        ldz = ldc if ldz is None else ldz
        fz = _flat(z)
        zidx = (z_offset + b0 * z_sb0 + b1 * z_sb1 + m * ldz + n).expand(nb1, nb0, M, N)
        assert zidx.min() >= 0 and zidx.max() < fz.numel(), "pre-activation view reaches outside its buffer"
        zz = fz[zidx].float()
        gz, dgz = (_gelu(zz), _gelu_grad(zz)) if act == 1 else (zz, torch.ones_like(zz))
        cm = _flat(colmask)[b0 * vec_sb0 + n] if colmask is not None else 1.0
        vidx = (b0 * vec_sb0 + n).expand(nb1, nb0, M, N).reshape(-1)
        out = v * cm * dgz
        if dbias is not None:
            _flat(dbias).index_add_(0, vidx, out.reshape(-1))
        if dmask is not None:
            _flat(dmask).index_add_(0, vidx, (v * gz).reshape(-1))
        fc = _flat(c)
        assert cidx.min() >= 0 and cidx.max() < fc.numel(), "GEMM output view reaches outside its buffer"
        fc[cidx.reshape(-1)] = out.reshape(-1).to(c.dtype)
        return c
    if c2 is not None:
        f2 = _flat(c2)
        assert cidx.max() < f2.numel()
        f2[cidx.reshape(-1)] = v.reshape(-1).to(c2.dtype)
    if act == 1:
        v = _gelu(v)
    if colmask is not None:
        v = v * _flat(colmask)[b0 * vec_sb0 + n]
    if scalar_mask is not None:
        v = v * _flat(scalar_mask)[b0 * scalar_sb0]
    if residual is not None:
        ridx = r_offset + b0 * r_sb0 + b1 * r_sb1 + m * ldr + n
        fr = _flat(residual)
        assert ridx.max() < fr.numel()
        v = v + fr[ridx].float()
    fc = _flat(c)
    assert cidx.min() >= 0 and cidx.max() < fc.numel(), "GEMM output view reaches outside its buffer"
    if atomic or accumulate:
        fc.index_add_(0, cidx.reshape(-1), v.reshape(-1).to(c.dtype))
    else:
        fc[cidx.reshape(-1)] = v.reshape(-1).to(c.dtype)
    return c


# ------------------------------------------------------------------------------------------------ small utilities
def memset(t, value=0):
    assert value == 0
    t.zero_()
    return t


def memset2d(t, offset_elems, pitch_elems, width_elems, height):
    f = _flat(t)
    for r in range(height):
        f[offset_elems + r * pitch_elems: offset_elems + r * pitch_elems + width_elems] = 0


def zeros(shape, dtype, device):
    return torch.zeros(shape, dtype=dtype, device=device)


def cast(src, dst):
    dst.copy_(src.reshape(dst.shape))
    return dst


def vec_mul(a, b, out, alpha=1.0, beta=0.0):
    v = a * (b if b is not None else 1.0) * alpha
    if beta != 0.0:
        v = v + beta * out
    out.copy_(v)
    return out


def transpose(src, dst, B, R, C):
    dst.view(B, C, R).copy_(src.view(B, R, C).transpose(1, 2))
    return dst


def axpy(dst, src, n=None):
    n = dst.numel() if n is None else n
    fd, fs = _flat(dst), _flat(src)
    fd[:n] += fs[:n]
    return dst


def group_mask_bwd(dy, z, mask, dout, dmask, rows, C, group):
    d = dy.reshape(rows, C).float()
    dmask += (d * z.reshape(rows, C).float()).sum(0).view(C // group, group).sum(1)
    dout.reshape(rows, C).copy_(d * mask.repeat_interleave(group))


def wave_norm(wave, out, lengths, B, T0, eps=1e-5):
    for b in range(B):
        n = int(lengths[b]) if lengths is not None else T0
        out[b].zero_()
        out[b, :n] = torch.nn.functional.layer_norm(wave[b, :n], (n,), eps=eps)
    return out


# ------------------------------------------------------------------------------------------------ norms / epilogue
def layernorm_fwd(x, w, b, y, rows, D, eps, mean=None, rstd=None, y_pre=None, colmask=None, act=0, drop_p=0.0, seed=0,
                  d_valid=0):
    assert drop_p == 0.0
    Dv = d_valid or D
    valid = (torch.arange(D) < Dv).float()
    xf = x.reshape(rows, D).float() * valid
    mu = xf.sum(1, keepdim=True) / Dv
    var = (((xf - mu) * valid) ** 2).sum(1, keepdim=True) / Dv
    rs = torch.rsqrt(var + eps)
    o = ((xf - mu) * rs * w + b) * valid
    if mean is not None:
        mean.copy_(mu.view(-1))
    if rstd is not None:
        rstd.copy_(rs.view(-1))
    if y_pre is not None:
        y_pre.reshape(rows, D).copy_(o)
    if act == 1:
        o = _gelu(o)
    if colmask is not None:
        o = o * colmask
    y.reshape(rows, D).copy_(o)
    return y


def _scatter_rows(dst, src, rows, C, rows_per_batch, out_bs, out_row_off):
    """rows of `src` (rows, C) -> dst[b * out_bs + (t + out_row_off) * C + c]"""
    if rows_per_batch <= 0:
        rows_per_batch, out_bs, out_row_off = rows, 0, 0
    f = _flat(dst)
    r = torch.arange(rows)
    b = r // rows_per_batch
    t = r - b * rows_per_batch
    idx = (b * out_bs + (t + out_row_off) * C).view(-1, 1) + torch.arange(C).view(1, -1)
    assert idx.max() < f.numel()
    f[idx.reshape(-1)] = src.reshape(-1).to(dst.dtype)


def layernorm_bwd(dy, x, w, b, mean, rstd, dx, dw, db, rows, D, dy2=None, colmask=None, dmask=None, act=0, drop_p=0.0,
                  seed=0, rows_per_batch=0, out_bs=0, out_row_off=0, d_valid=0):
    assert drop_p == 0.0
    Dv = d_valid or D
    valid = (torch.arange(D) < Dv).float()
    d = _flat(dy)[: rows * D].view(rows, D).float() * valid
    if dy2 is not None:
        d = d + _flat(dy2)[: rows * D].view(rows, D).float() * valid
    xh = (_flat(x)[: rows * D].view(rows, D).float() - mean.view(-1, 1)) * rstd.view(-1, 1) * valid
    if act == 1 or colmask is not None:
        u = xh * w + b
        m = colmask if colmask is not None else 1.0
        if act == 1:
            if dmask is not None:
                dmask += (d * _gelu(u)).sum(0)
            d = d * m * _gelu_grad(u)
        else:
            if dmask is not None:
                dmask += (d * u).sum(0)
            d = d * m
    if dw is not None:
        dw += (d * xh).sum(0)
    if db is not None:
        db += d.sum(0)
    g = d * w * valid
    s1 = g.sum(1, keepdim=True) / Dv
    s2 = (g * xh).sum(1, keepdim=True) / Dv
    res = rstd.view(-1, 1) * (g - s1 - xh * s2) * valid
    _scatter_rows(dx, res, rows, D, rows_per_batch, out_bs, out_row_off)
    return dx


def epilogue_bwd(dy, rows, C, z=None, colmask=None, scalar=None, dout=None, dmask=None, dbias=None, dscalar=None, act=0,
                 drop_p=0.0, seed=0, rows_per_batch=0, out_bs=0, out_row_off=0, in_ld=0, z_ld=0):
    assert drop_p == 0.0
    in_ld = in_ld or C
    z_ld = z_ld or C
    assert not ((act == 1 or dmask is not None or dscalar is not None) and z is None)
    d = torch.as_strided(_flat(dy), (rows, C), (in_ld, 1)).float()
    zz = torch.as_strided(_flat(z), (rows, C), (z_ld, 1)).float() if z is not None else None
    cm = colmask if colmask is not None else torch.ones(C)
    sc = scalar.reshape(-1)[0] if scalar is not None else 1.0
    az = (_gelu(zz) if act == 1 else zz) if zz is not None else torch.zeros(rows, C)
    if dscalar is not None:
        dscalar += (d * az * cm).sum()
    dm = d * sc
    if dmask is not None:
        dmask += (dm * az).sum(0)
    dv = dm * cm
    if act == 1:
        dv = dv * _gelu_grad(zz)
    if dbias is not None:
        dbias += dv.sum(0)
    if dout is not None:
        _scatter_rows(dout, dv, rows, C, rows_per_batch, out_bs, out_row_off)


def pad_rows_zero(x, lengths, B, T, D):
    xv = x.view(B, T, D)
    for b in range(B):
        n = max(int(lengths[b]), 0)
        xv[b, n:] = 0


def conv_out_lengths(lengths, out, ksizes, strides):
    L = lengths.clone()
    for k, s in zip(ksizes, strides):
        L = torch.clamp(torch.div(L - k, s, rounding_mode="floor") + 1, min=0)
    out.copy_(L)
    return out


def posconv_pad(x, xp, B, T, G_, Cg, Cgp, front, pad):
    xv = x.reshape(B, T, G_, Cg)
    Wd = G_ * Cgp
    f = _flat(xp)
    for b in range(B):
        r0 = front + b * (T + pad)
        blk = f[r0 * Wd:(r0 + T) * Wd].view(T, G_, Cgp)
        blk.zero_()
        blk[:, :, :Cg] = xv[b]


def posconv_tc(xp, wp, bias, residual, out, c2, B, T, D, G_, Cg, Cgp, K, rows_b, row0, act):
    xf = _flat(xp).view(-1, G_, Cgp).float()
    w = wp.reshape(G_, Cg, K, Cgp).float()
    ov = out.view(B, T, D)
    for b in range(B):
        idx = torch.arange(T)[:, None] + torch.arange(K)[None, :] + (row0 + b * rows_b)
        z = torch.einsum("tkgc,gnkc->tgn", xf[idx], w).reshape(T, D)
        if bias is not None:
            z = z + bias.float()
        if c2 is not None:
            c2.view(B, T, D)[b] = z.to(c2.dtype)
        y = torch.nn.functional.gelu(z) if act == 1 else z
        if residual is not None:
            y = y + residual.view(B, T, D)[b].float()
        ov[b] = y.to(out.dtype)


# ------------------------------------------------------------------------------------------------ conv layer 0
def _conv0_z(wave, w, bias, B, C, k, stride, T1):
    z = torch.nn.functional.conv1d(wave[:, None, :].float(), w.view(C, 1, k), bias, stride=stride)  # (B, C, T)
    return z[:, :, :T1]


def conv0_stats(wave, w, bias, moments, mean, rstd, B, C, k, stride, T1, eps):
    z = _conv0_z(wave, w, bias, B, C, k, stride, T1).double()
    mu = z.mean(2)
    var = ((z - mu[:, :, None]) ** 2).mean(2)
    mean.copy_(mu.reshape(-1).float())
    rstd.copy_((1.0 / torch.sqrt(var + eps)).reshape(-1).float())


def conv0_fwd(wave, w, bias, gamma, beta, mean, rstd, mask, y, B, C, k, stride, T1, norm, act):
    z = _conv0_z(wave, w, bias, B, C, k, stride, T1)
    if norm:
        z = (z - mean.view(B, C, 1)) * rstd.view(B, C, 1) * gamma.view(1, C, 1) + beta.view(1, C, 1)
    if act:
        z = _gelu(z)
    if mask is not None:
        z = z * mask.view(1, C, 1)
    y.view(B, T1, C).copy_(z.transpose(1, 2))


def conv0_bwd(wave, w, bias, gamma, beta, mean, rstd, moments, mask, dy, scratch, dW, dbias, dgamma, dbeta, dmask, B, C, k,
              stride, T1, norm, act):
    z = _conv0_z(wave, w, bias, B, C, k, stride, T1)  # (B, C, T)
    d = _flat(dy)[: B * T1 * C].view(B, T1, C).float().transpose(1, 2)
    if norm:
        yn = (z - mean.view(B, C, 1)) * rstd.view(B, C, 1)
        u = yn * gamma.view(1, C, 1) + beta.view(1, C, 1)
    else:
        yn, u = z, z
    au = _gelu(u) if act else u
    if dmask is not None:
        dmask += (d * au).sum((0, 2))
    du = d * (mask.view(1, C, 1) if mask is not None else 1.0) * (_gelu_grad(u) if act else 1.0)
    if norm:
        Am = du.mean(2, keepdim=True)
        Bm = (du * yn).mean(2, keepdim=True)
        dz = rstd.view(B, C, 1) * gamma.view(1, C, 1) * (du - Am - yn * Bm)
        if dgamma is not None:
            dgamma += (du * yn).sum((0, 2))
        if dbeta is not None:
            dbeta += du.sum((0, 2))
    else:
        dz = du
    # dW[c, q] = sum_{b,t} dz[b,c,t] * x[b, stride t + q]
    xw = wave.float().unfold(1, k, stride)[:, :T1]  # (B, T1, k)
    dW += torch.einsum("bct,btq->cq", dz, xw).view(dW.shape)
    if dbias is not None:
        dbias += dz.sum((0, 2))


# ------------------------------------------------------------------------------------------------ attention
def _bias_term(wavlm, B, H, T):
    rt, emb, gate, hidx, htot = wavlm
    i = torch.arange(T).view(-1, 1)
    j = torch.arange(T).view(1, -1)
    bucket = rt.long()[j - i + T - 1]  # (T, T)
    heads = hidx.long() if hidx is not None else torch.arange(H)
    bias = emb[bucket][:, :, heads].permute(2, 0, 1).unsqueeze(0).expand(B, H, T, T)  # (B, H, T, T)
    if gate is not None:
        bias = bias * gate.view(B, htot, T)[:, heads].unsqueeze(-1)
    return bias, bucket, heads


def softmax_fwd(S, lds, P, Pd, ldp, lengths, B, H, T, drop_p=0.0, seed=0, wavlm=None):
    assert drop_p == 0.0
    s = torch.as_strided(_flat(S), (B, H, T, T), (H * T * lds, T * lds, lds, 1)).float()
    if wavlm is not None:
        s = s + _bias_term(wavlm, B, H, T)[0]
    if lengths is not None:
        pad = torch.arange(T).view(1, -1) >= lengths.view(-1, 1)
        s = s + (-10000.0) * pad.view(B, 1, 1, T).float()
    p = torch.softmax(s, -1)
    for dst in (P, Pd):
        if dst is not None:
            d = torch.as_strided(_flat(dst), (B, H, T, ldp), (H * T * ldp, T * ldp, ldp, 1))
            d.zero_()
            d[..., :T] = p


def softmax_bwd(P, dPd, lds, dS, ldp, B, H, T, drop_p=0.0, seed=0, wavlm=None, dgate=None, demb=None, num_buckets=0):
    assert drop_p == 0.0
    p = torch.as_strided(_flat(P), (B, H, T, T), (H * T * ldp, T * ldp, ldp, 1)).float()
    d = torch.as_strided(_flat(dPd), (B, H, T, T), (H * T * lds, T * lds, lds, 1)).float()
    ds = p * (d - (d * p).sum(-1, keepdim=True))
    out = torch.as_strided(_flat(dS), (B, H, T, ldp), (H * T * ldp, T * ldp, ldp, 1))
    out.zero_()
    out[..., :T] = ds
    if wavlm is not None:
        rt, emb, gate, hidx, htot = wavlm
        _, bucket, heads = _bias_term(wavlm, B, H, T)
        embv = emb[bucket][:, :, heads].permute(2, 0, 1)  # (H, T, T)
        if dgate is not None:
            dgate.view(B, htot, T)[:, heads] = (ds * embv.unsqueeze(0)).sum(-1)
        if demb is not None:
            g = gate.view(B, htot, T)[:, heads].unsqueeze(-1) if gate is not None else 1.0
            contrib = (ds * g).sum(0)  # (H, T, T)
            for hi, ht in enumerate(heads.tolist()):
                demb[:, ht].index_add_(0, bucket.reshape(-1), contrib[hi].reshape(-1))


def wavlm_gate_fwd(x, Wg, bg, cst, gate, ga, gb, B, T, Htot, Dh):
    xv = x.reshape(B, T, Htot, Dh).float()
    g8 = torch.einsum("bthd,od->btho", xv, Wg) + bg
    a = torch.sigmoid(g8[..., :4].sum(-1))
    b = torch.sigmoid(g8[..., 4:].sum(-1))
    c = cst.reshape(Htot)
    gt = a * (b * c - 1.0) + 2.0
    gate.view(B, Htot, T).copy_(gt.permute(0, 2, 1))
    if ga is not None:
        ga.view(B, Htot, T).copy_(a.permute(0, 2, 1))
    if gb is not None:
        gb.view(B, Htot, T).copy_(b.permute(0, 2, 1))


def wavlm_gate_bwd(x, Wg, cst, ga, gb, dgate, dx, dWg, dbg, dcst, B, T, Htot, Dh):
    xv = x.reshape(B, T, Htot, Dh).float()
    a = ga.view(B, Htot, T).permute(0, 2, 1)
    b = gb.view(B, Htot, T).permute(0, 2, 1)
    dg = dgate.view(B, Htot, T).permute(0, 2, 1)
    c = cst.reshape(Htot)
    dcst.view(-1).add_((dg * a * b).sum((0, 1)))
    dpa = dg * (b * c - 1.0) * a * (1 - a)
    dpb = dg * a * c * b * (1 - b)
    dp = torch.cat([dpa.unsqueeze(-1).expand(-1, -1, -1, 4), dpb.unsqueeze(-1).expand(-1, -1, -1, 4)], -1)  # (B,T,H,8)
    dbg += dp.sum((0, 1, 2))
    dWg += torch.einsum("btho,bthd->od", dp, xv)
    if dx is not None:
        dxv = dx.view(B, T, Htot, Dh)
        dxv += torch.einsum("btho,od->bthd", dp, Wg).to(dx.dtype)


def _attn_core(qkv, B, H, T, scale, lengths, wavlm):
    Dh = 64
    inner = H * Dh
    x = qkv.reshape(B, T, 3, H, Dh).float()
    k, v, q = x[:, :, 0].permute(0, 2, 1, 3), x[:, :, 1].permute(0, 2, 1, 3), x[:, :, 2].permute(0, 2, 1, 3)
    s = scale * q @ k.transpose(-1, -2)
    if wavlm is not None:
        s = s + _bias_term(wavlm, B, H, T)[0]
    if lengths is not None:
        pad = torch.arange(T).view(1, -1) >= lengths.view(-1, 1)
        s = s + (-10000.0) * pad.view(B, 1, 1, T).float()
    return q, k, v, s


def attn_fwd(qkv, ctx, ctx_un, lse, B, H, T, scale, lengths=None, drop_p=0.0, seed=0, wavlm=None, head_mask=None):
    assert drop_p == 0.0 and qkv.dtype == torch.bfloat16
    q, k, v, s = _attn_core(qkv, B, H, T, scale, lengths, wavlm)
    p = torch.softmax(s, -1)
    o = (p @ v).permute(0, 2, 1, 3).reshape(B * T, H * 64)
    if lse is not None:
        lse.view(B, H, T).copy_(torch.logsumexp(s, -1))
    if ctx_un is not None:
        ctx_un.view(B * T, -1).copy_(o)
    if head_mask is not None:
        o = o * head_mask.repeat_interleave(64)
    ctx.view(B * T, -1).copy_(o)


def attn_bwd(qkv, dctx, o_un, lse, delta, dq32, dqkv, dmask_h, B, H, T, scale, lengths=None, drop_p=0.0, seed=0,
             wavlm=None, head_mask=None, dgate=None, demb=None, num_buckets=0):
    assert drop_p == 0.0
    q, k, v, s = _attn_core(qkv, B, H, T, scale, lengths, wavlm)
    p = torch.softmax(s, -1)
    d = dctx.reshape(B, T, H, 64).float().permute(0, 2, 1, 3)  # gradient wrt the masked context
    ou = o_un.reshape(B, T, H, 64).float().permute(0, 2, 1, 3)
    dl = (d * ou).sum(-1)  # (B, H, T)
    delta.view(B, H, T).copy_(dl)
    if dmask_h is not None:
        dmask_h += dl.sum((0, 2))
    hm = head_mask.view(1, H, 1, 1) if head_mask is not None else 1.0
    do = d * hm
    dv = p.transpose(-1, -2) @ do
    dp = do @ v.transpose(-1, -2)
    ds = p * (dp - (dp * p).sum(-1, keepdim=True))
    dq = scale * ds @ k
    dk = scale * ds.transpose(-1, -2) @ q
    out = dqkv.view(B, T, 3, H, 64)
    out[:, :, 0] = dk.permute(0, 2, 1, 3).to(dqkv.dtype)
    out[:, :, 1] = dv.permute(0, 2, 1, 3).to(dqkv.dtype)
    out[:, :, 2] = dq.permute(0, 2, 1, 3).to(dqkv.dtype)
    if wavlm is not None:
        rt, emb, gate, hidx, htot = wavlm
        _, bucket, heads = _bias_term(wavlm, B, H, T)
        embv = emb[bucket][:, :, heads].permute(2, 0, 1)
        if dgate is not None:
            dgate.view(B, htot, T)[:, heads] += (ds * embv.unsqueeze(0)).sum(-1)
        if demb is not None:
            g = gate.view(B, htot, T)[:, heads].unsqueeze(-1) if gate is not None else 1.0
            contrib = (ds * g).sum(0)
            for hi, ht in enumerate(heads.tolist()):
                demb[:, ht].index_add_(0, bucket.reshape(-1), contrib[hi].reshape(-1))


# ------------------------------------------------------------------------------------------------ loss
def distill_loss_fwd(a, b, rowstat, acc, out, rows, D, w2, w1, wc, cos_type, first, last, total_rows):
    x, y = a.reshape(rows, D).float(), b.reshape(rows, D).float()
    if first:
        acc.zero_()
    d = x - y
    dot = (x * y).sum(1)
    na, nb = x.norm(dim=1), y.norm(dim=1)
    rowstat.view(-1)[: rows * 3].view(rows, 3).copy_(torch.stack([dot, na, nb], 1))
    c = dot / (na.clamp_min(1e-8) * nb.clamp_min(1e-8))
    cv = c if cos_type == 0 else -torch.log1p(torch.exp(-c))
    acc[0] += (d * d).sum()
    acc[1] += d.abs().sum()
    acc[2] += cv.sum()
    if last:
        ne, nr = float(total_rows) * D, float(total_rows)
        mse = acc[0] / ne if w2 != 0 else 0.0
        l1 = acc[1] / ne if w1 != 0 else 0.0
        lc = -acc[2] / nr if wc != 0 else 0.0
        out[0] = w2 * mse + w1 * l1 + wc * lc
        out[1], out[2], out[3] = mse, l1, lc


def distill_loss_bwd(a, b, rowstat, upstream, da, rows, total_rows, D, w2, w1, wc, cos_type):
    x, y = a.reshape(rows, D).float(), b.reshape(rows, D).float()
    g = upstream.reshape(-1)[0] if upstream is not None else 1.0
    rs = rowstat.view(-1)[: rows * 3].view(rows, 3)
    dot, na, nb = rs[:, 0:1], rs[:, 1:2], rs[:, 2:3]
    eps = 1e-8
    nac, nbc = na.clamp_min(eps), nb.clamp_min(eps)
    c = dot / (nac * nbc)
    k = torch.full_like(c, -wc / float(total_rows))
    if cos_type != 0:
        k = k / (1.0 + torch.exp(c))
    kb = k / (nac * nbc)
    ka = torch.where(na > eps, k * c / (na * na), torch.zeros_like(c))
    ne = float(total_rows) * D
    d = x - y
    gr = 2.0 * w2 / ne * d + w1 / ne * torch.sign(d) + kb * y - ka * x
    da.reshape(rows, D).copy_(g * gr)


# ------------------------------------------------------------------------------------------------ weight prep
def conv_weight_prep(w, wf, wd, Cout, Cin, k, perm):
    wv = w.view(Cout, Cin, k)
    if wf is not None:
        wf.view(Cout, k, Cin).copy_(wv.permute(0, 2, 1))
    if wd is not None:
        wdv = wd.view(k, Cin, Cout)
        for tap in range(k):
            wdv[perm[tap]].copy_(wv[:, :, tap].t())


def conv_wgrad_unpermute(src, dw, Cout, Cin, k):
    dw.view(Cout, Cin, k).add_(src.view(Cout, k, Cin).permute(0, 2, 1))


def posconv_weight_prep(v, g, norm, wn, wp, wdp, D, G_, Cg, Cgp, K):
    nrm = torch.linalg.vector_norm(v, dim=(0, 1))
    norm.copy_(nrm)
    w = g.view(1, 1, K) * v / nrm.view(1, 1, K)  # (D, Cg, K)
    if wn is not None:
        wn.view(D, Cg, K).copy_(w)
    wg = w.view(G_, Cg, Cg, K)  # (g, co, ci, tap)
    if wp is not None:
        wpv = wp.view(G_, Cg, K, Cgp)
        wpv.zero_()
        wpv[..., :Cg] = wg.permute(0, 1, 3, 2)
    if wdp is not None:
        wdv = wdp.view(G_, Cg, K, Cgp)
        wdv.zero_()
        wdv[..., :Cg] = wg.permute(0, 2, 3, 1).flip(2)  # (g, ci, K-1-tap, co)


def posconv_weight_bwd(dwp, v, g, norm, scratch, dv, dg, D, G_, Cg, Cop, Cgp, K):
    dw = dwp.view(G_, Cop, K, Cgp)[:, :Cg, :, :Cg].permute(0, 1, 3, 2).reshape(D, Cg, K)  # (d, ci, tap)
    s = (dw * v).sum((0, 1))
    dg.view(-1).add_(s / norm)
    dv.add_(g.view(1, 1, K) / norm.view(1, 1, K) * (dw - v * (s / (norm * norm)).view(1, 1, K)))


# ------------------------------------------------------------------------------------------------ optimizer
def sumsq(x, out):
    out += (x.double() ** 2).sum()


def counter_inc(c):
    c += 1


def adamw(p, g, m, v, lr_dev, beta1, beta2, eps, wd, step_dev, gnorm_sq, max_norm, p_bf16, zero_grad):
    lr = float(lr_dev.reshape(-1)[0])
    step = int(step_dev.reshape(-1)[0])
    clip = 1.0
    if gnorm_sq is not None and max_norm > 0:
        clip = min(1.0, max_norm / (math.sqrt(float(gnorm_sq.reshape(-1)[0])) + 1e-6))
    gr = g * clip
    p.mul_(1.0 - lr * wd)
    m.mul_(beta1).add_(gr, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(gr, gr, value=1 - beta2)
    bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
    denom = v.sqrt() / math.sqrt(bc2) + eps
    p.addcdiv_(m, denom, value=-lr / bc1)
    if p_bf16 is not None:
        p_bf16.view(-1)[: p.numel()].copy_(p.view(-1))
    if zero_grad:
        g.zero_()


def scale_(x, a):
    x.mul_(a)


# ------------------------------------------------------------------------------------------------ HardConcrete
def hardconcrete_fwd(la, u_in, u_out, mask, n, beta, l, r, eps, seed):
    la = la.view(-1)[:n]
    if u_in is None:
        u = torch.empty(n).uniform_(eps, 1 - eps)
    else:
        u = u_in.view(-1)[:n]
    if u_out is not None:
        u_out.view(-1)[:n].copy_(u)
    s = torch.sigmoid((torch.log(u / (1 - u)) + la) / beta)
    mask.view(-1)[:n].copy_((s * (r - l) + l).clamp(0.0, 1.0))


def hardconcrete_bwd(la, u, dmask, dla, n, beta, l, r):
    la = la.view(-1)[:n]
    uu = u.view(-1)[:n]
    s = torch.sigmoid((torch.log(uu / (1 - uu)) + la) / beta)
    m = s * (r - l) + l
    g = torch.where((m > 0) & (m < 1), dmask.view(-1)[:n] * (r - l) * s * (1 - s) / beta, torch.zeros(n))
    dla.view(-1)[:n].add_(g)


def l0_norms(la_base, seg_off, seg_len, nseg, bias, norms):
    la = _flat(la_base)
    for s in range(nseg):
        o, n = int(seg_off[s]), int(seg_len[s])
        norms[s] = torch.sigmoid(la[o:o + n] + bias).sum()


def num_params(prog, norms, nseg, out_n, dnorm):
    with torch.enable_grad():
        _num_params(prog, norms, nseg, out_n, dnorm)


def _num_params(prog, norms, nseg, out_n, dnorm):
    """Same program walk as csrc/hardconcrete.cu:num_params_kernel, written with autograd for the derivative."""
    nr = norms.detach().clone().double().requires_grad_(True)
    pr = [int(x) for x in prog.tolist()]
    p = 0
    n_conv = pr[p]; p += 1
    total = torch.zeros((), dtype=torch.float64)
    in_ch = torch.ones((), dtype=torch.float64)
    for _ in range(n_conv):
        seg, k, hb, hn, so = pr[p:p + 5]; p += 5
        out = nr[seg] if seg >= 0 else torch.tensor(float(so), dtype=torch.float64)
        total = total + in_ch * out * k + (out if hb else 0.0) + (2.0 * out if hn else 0.0)
        in_ch = out
    proj_out, static_extra, D, n_layers = pr[p:p + 4]; p += 4
    total = total + in_ch + in_ch * 2.0 + (in_ch + 1.0) * proj_out + static_extra
    for _ in range(n_layers):
        att, hseg, hst, hd, ags, ff, fseg, fst, fgs = pr[p:p + 9]; p += 9
        total = total + 4.0 * D
        if att:
            h = nr[hseg] if hseg >= 0 else float(hst)
            a = (D + 1.0) * h * hd * 3.0 + (h * hd + 1.0) * D
            total = total + (a * nr[ags] if ags >= 0 else a)
        if ff:
            f = nr[fseg] if fseg >= 0 else float(fst)
            a = (D + 1.0) * f + (f + 1.0) * D
            total = total + (a * nr[fgs] if fgs >= 0 else a)
    out_n[0] = float(total.detach())
    if nseg:
        (gr,) = torch.autograd.grad(total, nr, allow_unused=True)
        dnorm.copy_(gr.float() if gr is not None else torch.zeros(nseg))


def lagrangian(n_hat, lambdas, n0, target, out):
    s = 1.0 - float(n_hat.reshape(-1)[0]) / n0
    t = float(target.reshape(-1)[0])
    l1, l2 = float(lambdas.reshape(-1)[0]), float(lambdas.reshape(-1)[1])
    d = s - t
    out[0] = l1 * d + l2 * d * d
    out[1] = s
    out[2] = -(l1 + 2.0 * l2 * d) / n0
    out[3] = d
    out[4] = d * d


def l0_norms_bwd(la_base, dla_base, seg_off, seg_len, nseg, bias, dnorm, upstream, coef):
    la, dla = _flat(la_base), _flat(dla_base)
    for s in range(nseg):
        o, n = int(seg_off[s]), int(seg_len[s])
        g = float(dnorm[s])
        if upstream is not None:
            g *= float(upstream.reshape(-1)[0])
        if coef is not None:
            g *= float(coef.reshape(-1)[0])
        sg = torch.sigmoid(la[o:o + n] + bias)
        dla[o:o + n] += g * sg * (1 - sg)


def collate_crop(pool, start, length, out, lengths_out, B, T):
    out.zero_()
    for b in range(B):
        n = min(int(length[b]), T)
        src = pool[int(start[b]): int(start[b]) + n]
        out[b, :n] = src.float() / 32768.0 if pool.dtype == torch.int16 else src
        if lengths_out is not None:
            lengths_out[b] = n
    return out


_EPOCH = [0]


def rng_set(value):
    _EPOCH[0] = int(value)


def rng_advance():
    _EPOCH[0] += 1


def rng_get(device):
    return _EPOCH[0]


_NAMES = ["memset", "memset2d", "zeros", "cast", "vec_mul", "transpose", "axpy", "group_mask_bwd", "wave_norm",
          "layernorm_fwd", "layernorm_bwd", "epilogue_bwd", "pad_rows_zero", "conv_out_lengths", "posconv_pad", "posconv_tc", "conv0_stats", "conv0_fwd",
          "conv0_bwd", "softmax_fwd", "softmax_bwd", "attn_fwd", "attn_bwd", "wavlm_gate_fwd", "wavlm_gate_bwd", "distill_loss_fwd",
          "distill_loss_bwd", "conv_weight_prep", "conv_wgrad_unpermute", "posconv_weight_prep", "posconv_weight_bwd",
          "sumsq", "counter_inc", "adamw", "scale_", "hardconcrete_fwd", "hardconcrete_bwd", "l0_norms", "num_params",
          "lagrangian", "l0_norms_bwd", "rng_set", "rng_advance", "rng_get", "collate_crop"]


def install(monkeypatch):
    """Route every kernel wrapper to its CPU restatement for the duration of a test."""
    g = globals()
    for n in _NAMES:
        assert hasattr(ops, n), n
        monkeypatch.setattr(ops, n, g[n])
    monkeypatch.setattr(G, "gemm", gemm)
    monkeypatch.setattr(engine, "_require_cuda", lambda t, what: None)
    monkeypatch.setattr(engine, "_is_native", lambda t: True)
    monkeypatch.setattr(engine, "_PLANS", {})
    monkeypatch.setattr(engine, "_SHADOWS", {})
