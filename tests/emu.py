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
         accumulate=False, atomic=False, ksplit=1, drop_p=0.0, seed=0, block_n=0, vec_sb0=0, scalar_sb0=0):
    assert drop_p == 0.0, "the emulator runs the dropout-free parity configuration"
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
