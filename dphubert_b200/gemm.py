"""Python-side descriptors for the C-ABI GEMMs (`dph_gemm_bf16`, `dph_gemm_f32`).

An operand is a strided *view* described by plain integers; strided convolutions, grouped convolutions and
per-head attention batches are all expressed as views of the activation buffers, never as copies.
"""
import ctypes

import torch

from . import _lib
from ._lib import GemmDesc, GemmF32Desc, Operand, OperandF32

BF16 = 0
F32 = 1

# When set to a list, every tensor-core GEMM launch appends (descriptor, algorithmic flops): bench.py replays exactly
# these launches to time the dominant kernel in isolation for the roofline figure.
TRACE = None


def dtype_code(t):
    if t.dtype == torch.bfloat16:
        return BF16
    if t.dtype == torch.float32:
        return F32
    raise _lib.DphError("unsupported dtype %s" % t.dtype)


class View:
    """Dtype-agnostic operand view.  element(row, k, b0, b1) =
    base[offset + b0*sb0 + b1*sb1 + row*rs + (k // kin)*kxs + (k % kin)*cs].
    For the bf16 tensor-core path either cs == 1 (K-major) or rs == 1 (MN-major, kin == K) must hold."""

    __slots__ = ("t", "offset", "rows", "K", "rs", "cs", "kin", "kxs", "sb0", "sb1")

    def __init__(self, t, rows, K, rs, cs=1, kin=None, kxs=0, sb0=0, sb1=0, offset=0):
        self.t, self.rows, self.K, self.rs, self.cs = t, rows, K, rs, cs
        self.kin = K if kin is None else kin
        self.kxs, self.sb0, self.sb1, self.offset = kxs, sb0, sb1, offset


def mat(t):
    """2-D row-major tensor (rows, K) as a K-major view."""
    assert t.dim() == 2 and t.stride(1) == 1
    return View(t, t.shape[0], t.shape[1], t.stride(0))


def mat_t(t):
    """2-D row-major tensor (K, rows) used transposed: MN-major view (reduction index is the slow dim)."""
    assert t.dim() == 2 and t.stride(1) == 1
    return View(t, t.shape[1], t.shape[0], 1, cs=t.stride(0))


def _operand_bf16(v, nb0, nb1, o=None):
    o = Operand() if o is None else o
    o.ptr = v.t.data_ptr() + v.offset * 2
    d3 = nb0 if v.sb0 != 0 else 1
    d4 = nb1 if v.sb1 != 0 else 1
    if v.cs == 1:
        o.major_mn = 0
        kx = (v.K + v.kin - 1) // v.kin
        o.dims[:] = (v.kin, v.rows, kx, d3, d4)
        o.strides[:] = (v.rs, v.kxs if kx > 1 else 8, v.sb0 if d3 > 1 else 8, v.sb1 if d4 > 1 else 8)
    elif v.rs == 1 and v.kin == v.K:
        o.major_mn = 1
        o.dims[:] = (v.rows, v.K, d3, d4, 1)
        o.strides[:] = (v.cs, v.sb0 if d3 > 1 else 8, v.sb1 if d4 > 1 else 8, 8)
    else:
        raise _lib.DphError("bf16 GEMM operand must be K-major (cs==1) or MN-major (rs==1)")
    o.b0_mul = 1 if d3 > 1 else 0
    o.b1_mul = 1 if d4 > 1 else 0
    return o


def _operand_f32(v):
    o = OperandF32()
    o.ptr = v.t.data_ptr() + v.offset * 4
    o.rs, o.cs, o.kxs, o.kin, o.bs0, o.bs1 = v.rs, v.cs, v.kxs, v.kin, v.sb0, v.sb1
    return o


def gemm(a, b, c, M, N, K, *, ldc=None, nb0=1, nb1=1, c_sb0=0, c_sb1=0, c_offset=0, alpha=1.0, bias=None, act=0,
         colmask=None, scalar_mask=None, residual=None, ldr=None, r_sb0=0, r_sb1=0, r_offset=0, c2=None,
         accumulate=False, atomic=False, ksplit=1, drop_p=0.0, seed=0, block_n=0, vec_sb0=0, scalar_sb0=0,
         cta_group=0, act_bwd=False, z=None, ldz=None, z_sb0=0, z_sb1=0, z_offset=0, dbias=None, dmask=None,
         b_taps=0):
    """C[b][m, n] = epilogue(alpha * sum_k a(m,k) b(n,k)).  `a`, `b` are `View`s; dtype of `a.t` selects the path:
    bf16 -> tcgen05 tensor-core kernel, fp32 -> CUDA-core kernel.
    `b_taps` > 0 (bf16, MN-major `b` with rows = N / b_taps channels): column n reads channel n % rows of tap n // rows,
    `b.sb1` being the element offset between taps (overlapping view; grouped-conv weight gradient).
    `act_bwd` (bf16 only, see `act_bwd_ok`): the product is the gradient wrt dropout(colmask * act(z)); the epilogue
    writes acc * dropout * colmask * act'(z) and accumulates `dbias` / `dmask` column sums."""
    L = _lib.lib()
    if ldc is None:
        ldc = N
    if ldr is None:
        ldr = ldc
    if a.t.dtype == torch.bfloat16:
        assert b.t.dtype == torch.bfloat16
        d = GemmDesc()
        d.a = _operand_bf16(a, nb0, nb1)
        d.b = _operand_bf16(b, nb0, b_taps if b_taps else nb1)
        if b_taps:
            assert nb1 == 1 and b.rows * b_taps == N and b.sb1 != 0
            d.b_nsplit = b.rows
        d.M, d.N, d.K, d.nb0, d.nb1 = M, N, K, nb0, nb1
        d.block_n, d.ksplit, d.cta_group = block_n, ksplit, cta_group
        d.c = c.data_ptr() + c_offset * c.element_size()
        d.c2 = None if c2 is None else c2.data_ptr() + c_offset * c2.element_size()
        d.ldc, d.c_sb0, d.c_sb1 = ldc, c_sb0, c_sb1
        d.c_dtype = dtype_code(c)
        d.c_mode = 1 if atomic else (2 if accumulate else 0)
        d.alpha = alpha
        d.bias = None if bias is None else bias.data_ptr()
        d.act = act
        d.colmask = None if colmask is None else colmask.data_ptr()
        d.scalar_mask = None if scalar_mask is None else scalar_mask.data_ptr()
        if residual is not None:
            d.residual = residual.data_ptr() + r_offset * residual.element_size()
            d.r_dtype = dtype_code(residual)
        d.ldr, d.r_sb0, d.r_sb1 = ldr, r_sb0, r_sb1
        d.drop_p, d.seed = drop_p, seed
        d.vec_sb0, d.scalar_sb0 = vec_sb0, scalar_sb0
        if act_bwd:
            d.act_bwd = 1
            d.z = z.data_ptr() + z_offset * z.element_size()
            d.ldz, d.z_sb0, d.z_sb1 = (ldc if ldz is None else ldz), z_sb0, z_sb1
            d.dbias = None if dbias is None else dbias.data_ptr()
            d.dmask = None if dmask is None else dmask.data_ptr()
        if TRACE is not None:
            TRACE.append((d, 2.0 * M * N * K * nb0 * nb1))
        _lib.check(L.dph_gemm_bf16(ctypes.byref(d), _lib.stream_ptr()), "dph_gemm_bf16")
    else:
        assert a.t.dtype == torch.float32 and b.t.dtype == torch.float32 and c.dtype == torch.float32
        if act_bwd:
            raise _lib.DphError("the fused activation backward exists on the bf16 tensor-core path only")
        d = GemmF32Desc()
        d.a = _operand_f32(a)
        d.b = _operand_f32(b)
        d.M, d.N, d.K, d.nb0, d.nb1 = M, N, K, nb0, nb1
        d.c = c.data_ptr() + c_offset * 4
        d.c2 = None if c2 is None else c2.data_ptr() + c_offset * 4
        d.ldc, d.c_sb0, d.c_sb1 = ldc, c_sb0, c_sb1
        d.accumulate = 2 if atomic else (1 if accumulate else 0)
        d.alpha = alpha
        d.bias = None if bias is None else bias.data_ptr()
        d.act = act
        d.colmask = None if colmask is None else colmask.data_ptr()
        d.scalar_mask = None if scalar_mask is None else scalar_mask.data_ptr()
        d.residual = None if residual is None else residual.data_ptr() + r_offset * 4
        d.ldr, d.r_sb0, d.r_sb1 = ldr, r_sb0, r_sb1
        d.drop_p, d.seed = drop_p, seed
        d.vec_sb0, d.scalar_sb0 = vec_sb0, scalar_sb0
        _lib.check(L.dph_gemm_f32(ctypes.byref(d), _lib.stream_ptr()), "dph_gemm_f32")
    return c


def act_bwd_ok(c, z, N, ldc, ldz, strides=(), offsets=()):
    """Can `gemm(..., act_bwd=True)` take this launch? bf16 in / out, whole 32-column chunks, 32-byte aligned rows."""
    if c.dtype != torch.bfloat16 or z is None or z.dtype != torch.bfloat16 or N % 32:
        return False
    if ldc % 16 or ldz % 16 or any(int(s_) % 16 for s_ in strides) or any(int(o) % 16 for o in offsets):
        return False
    return c.data_ptr() % 32 == 0 and z.data_ptr() % 32 == 0


def replay(trace):
    """Re-issue traced tensor-core GEMM launches (same descriptors, same buffers)."""
    L = _lib.lib()
    sp = _lib.stream_ptr()
    for d, _ in trace:
        _lib.check(L.dph_gemm_bf16(ctypes.byref(d), sp), "dph_gemm_bf16")
