"""Thin typed wrappers over the C-ABI entry points (one Python function per `dph_*` symbol) plus the few
stand-alone autograd ops the module mirror needs (HardConcrete sampling, l0 norm).

Everything here takes CUDA tensors, launches on torch's current stream and never falls back to torch math.
"""
import ctypes

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr
from .gemm import dtype_code

_NULL = None


def _L():
    return _lib.lib()


# ----------------------------------------------------------------------------------------------------------------
# small utilities
# ----------------------------------------------------------------------------------------------------------------
def memset(t, value=0):
    check(_L().dph_memset(ptr(t), value, t.numel() * t.element_size(), stream_ptr()), "dph_memset")
    return t


def memset2d(t, offset_elems, pitch_elems, width_elems, height):
    es = t.element_size()
    p = ctypes.c_void_p(t.data_ptr() + offset_elems * es)
    check(_L().dph_memset2d(p, pitch_elems * es, 0, width_elems * es, height, stream_ptr()), "dph_memset2d")


def zeros(shape, dtype, device):
    return memset(torch.empty(shape, dtype=dtype, device=device))


def cast(src, dst):
    check(_L().dph_cast(ptr(src), dtype_code(src), ptr(dst), dtype_code(dst), src.numel(), stream_ptr()), "dph_cast")
    return dst


def vec_mul(a, b, out, alpha=1.0, beta=0.0):
    check(_L().dph_vec_mul(ptr(a), ptr(b), ptr(out), alpha, beta, a.numel(), stream_ptr()), "dph_vec_mul")
    return out


def transpose(src, dst, B, R, C):
    """(B, R, C) -> (B, C, R)"""
    check(_L().dph_transpose(ptr(src), dtype_code(src), ptr(dst), dtype_code(dst), B, R, C, stream_ptr()), "dph_transpose")
    return dst


def axpy(dst, src, n=None):
    """dst += src (same dtype)"""
    n = dst.numel() if n is None else n
    check(_L().dph_axpy(ptr(dst), ptr(src), n, dtype_code(dst), stream_ptr()), "dph_axpy")
    return dst


def group_mask_bwd(dy, z, mask, dout, dmask, rows, C, group):
    check(_L().dph_group_mask_bwd(ptr(dy), ptr(z), ptr(mask), ptr(dout), ptr(dmask), rows, C, group, dtype_code(dy),
                                  stream_ptr()), "dph_group_mask_bwd")


# ----------------------------------------------------------------------------------------------------------------
# norms / elementwise
# ----------------------------------------------------------------------------------------------------------------
def layernorm_fwd(x, w, b, y, rows, D, eps, mean=None, rstd=None, y_pre=None, colmask=None, act=0, drop_p=0.0, seed=0,
                  d_valid=0):
    check(_L().dph_layernorm_fwd(ptr(x), ptr(w), ptr(b), ptr(y), ptr(y_pre), ptr(mean), ptr(rstd), ptr(colmask), rows, D,
                                 d_valid, eps, act, drop_p, seed, dtype_code(x), stream_ptr()), "dph_layernorm_fwd")
    return y


def layernorm_bwd(dy, x, w, b, mean, rstd, dx, dw, db, rows, D, dy2=None, colmask=None, dmask=None, act=0, drop_p=0.0,
                  seed=0, rows_per_batch=0, out_bs=0, out_row_off=0, d_valid=0):
    check(_L().dph_layernorm_bwd(ptr(dy), ptr(dy2), ptr(x), ptr(w), ptr(b), ptr(mean), ptr(rstd), ptr(dx), ptr(dw),
                                 ptr(db), ptr(colmask), ptr(dmask), rows, D, d_valid, act, drop_p, seed, rows_per_batch,
                                 out_bs, out_row_off, dtype_code(x), stream_ptr()), "dph_layernorm_bwd")
    return dx


def epilogue_bwd(dy, rows, C, z=None, colmask=None, scalar=None, dout=None, dmask=None, dbias=None, dscalar=None, act=0,
                 drop_p=0.0, seed=0, rows_per_batch=0, out_bs=0, out_row_off=0, in_ld=0, z_ld=0):
    check(_L().dph_epilogue_bwd(ptr(dy), ptr(z), ptr(colmask), ptr(scalar), ptr(dout), ptr(dmask), ptr(dbias),
                                ptr(dscalar), rows, C, act, drop_p, seed, rows_per_batch, out_bs, out_row_off, in_ld,
                                z_ld, dtype_code(dy), stream_ptr()), "dph_epilogue_bwd")


def pad_rows_zero(x, lengths, B, T, D):
    check(_L().dph_pad_rows_zero(ptr(x), ptr(lengths), B, T, D, dtype_code(x), stream_ptr()), "dph_pad_rows_zero")


def conv_out_lengths(lengths, out, ksizes, strides):
    """Frame counts after the conv stack for all utterances in one launch (components.py:116-119)."""
    import ctypes
    n = len(ksizes)
    ks = (ctypes.c_int * max(n, 1))(*ksizes)
    ss = (ctypes.c_int * max(n, 1))(*strides)
    check(_L().dph_conv_out_lengths(ptr(lengths), ptr(out), lengths.numel(), n, ks, ss, stream_ptr()),
          "dph_conv_out_lengths")
    return out


def posconv_pad(x, xp, B, T, G, Cg, Cgp, front, pad):
    check(_L().dph_posconv_pad(ptr(x), ptr(xp), B, T, G, Cg, Cgp, front, pad, dtype_code(x), stream_ptr()),
          "dph_posconv_pad")


# ----------------------------------------------------------------------------------------------------------------
# conv0
# ----------------------------------------------------------------------------------------------------------------
def posconv_tc(xp, wp, bias, residual, out, c2, B, T, D, G, Cg, Cgp, K, rows_b, row0, act):
    """Tap-stationary tensor-core positional conv (bf16): out = [residual +] act(conv(xp, wp) + bias)."""
    check(_L().dph_posconv_tc(ptr(xp), xp.shape[0], ptr(wp), ptr(bias), ptr(residual), ptr(out), ptr(c2), B, T, D, G, Cg,
                              Cgp, K, rows_b, row0, act, stream_ptr()), "dph_posconv_tc")


def posconv_tc_ok(dt, Cg, Cgp, K):
    return dt == torch.bfloat16 and Cgp == 64 and Cg % 16 == 0 and 16 <= Cg <= 64 and K % 8 == 0 and 64 <= K <= 128


def conv0_stats(wave, w, bias, moments, mean, rstd, B, C, k, stride, T1, eps):
    check(_L().dph_conv0_stats(ptr(wave), wave.stride(0), ptr(w), ptr(bias), ptr(moments), ptr(mean), ptr(rstd), B, C, k,
                               stride, T1, eps, stream_ptr()), "dph_conv0_stats")


def conv0_fwd(wave, w, bias, gamma, beta, mean, rstd, mask, y, B, C, k, stride, T1, norm, act):
    check(_L().dph_conv0_fwd(ptr(wave), wave.stride(0), ptr(w), ptr(bias), ptr(gamma), ptr(beta), ptr(mean), ptr(rstd),
                             ptr(mask), ptr(y), B, C, k, stride, T1, norm, act, dtype_code(y), stream_ptr()),
          "dph_conv0_fwd")


def conv0_bwd(wave, w, bias, gamma, beta, mean, rstd, moments, mask, dy, scratch, dW, dbias, dgamma, dbeta, dmask, B, C, k,
              stride, T1, norm, act):
    """`moments`: the (B*65,) fp64 buffer conv0_stats filled in the forward pass (GroupNorm mode); `scratch`: 16*B*C."""
    check(_L().dph_conv0_bwd(ptr(wave), wave.stride(0), ptr(w), ptr(bias), ptr(gamma), ptr(beta), ptr(mean), ptr(rstd),
                             ptr(moments), ptr(mask), ptr(dy), ptr(scratch), ptr(dW), ptr(dbias), ptr(dgamma),
                             ptr(dbeta), ptr(dmask), B, C, k, stride, T1, norm, act, dtype_code(dy), stream_ptr()),
          "dph_conv0_bwd")


def conv0_bwd_static(wave, w, bias, gamma, beta, mean, rstd, moments, mask, dy, dW, dbias, dgamma, dbeta, dmask, B, C, k,
                     stride, T1, norm, act):
    """`conv0_bwd` with the scratch in the library's file-scope device array (no allocation; B * C <= 16384; calls
    that overlap on the GPU race on the shared array).  Not used by the engine: for callers that cannot allocate."""
    check(_L().dph_conv0_bwd_static(ptr(wave), wave.stride(0), ptr(w), ptr(bias), ptr(gamma), ptr(beta), ptr(mean),
                                    ptr(rstd), ptr(moments), ptr(mask), ptr(dy), ptr(dW), ptr(dbias), ptr(dgamma),
                                    ptr(dbeta), ptr(dmask), B, C, k, stride, T1, norm, act, dtype_code(dy), stream_ptr()),
          "dph_conv0_bwd_static")


def wave_norm(wave, out, lengths, B, T0, eps=1e-5):
    check(_L().dph_wave_norm(ptr(wave), wave.stride(0), ptr(out), out.stride(0), ptr(lengths), B, T0, eps, stream_ptr()),
          "dph_wave_norm")
    return out


def collate_crop(pool, start, length, out, lengths_out, B, T):
    """out[b, :] = pool[start[b] : start[b] + length[b]] zero-padded to T (audio_dataset.py:285-363 on the device)"""
    code = {torch.float32: 1, torch.int16: 2}[pool.dtype]
    check(_L().dph_collate_crop(ptr(pool), code, ptr(start), ptr(length), ptr(out), ptr(lengths_out), B, T, stream_ptr()),
          "dph_collate_crop")
    return out


# ----------------------------------------------------------------------------------------------------------------
# attention probabilities
# ----------------------------------------------------------------------------------------------------------------
def softmax_fwd(S, lds, P, Pd, ldp, lengths, B, H, T, drop_p=0.0, seed=0, wavlm=None):
    rt = emb = gate = hidx = None
    htot = H
    if wavlm is not None:
        rt, emb, gate, hidx, htot = wavlm
    check(_L().dph_softmax_fwd(ptr(S), lds, ptr(P), ptr(Pd), ldp, ptr(lengths), B, H, T, drop_p, seed, ptr(rt), ptr(emb),
                               ptr(gate), ptr(hidx), htot, dtype_code(P), stream_ptr()), "dph_softmax_fwd")


def softmax_bwd(P, dPd, lds, dS, ldp, B, H, T, drop_p=0.0, seed=0, wavlm=None, dgate=None, demb=None, num_buckets=0):
    rt = emb = gate = hidx = None
    htot = H
    if wavlm is not None:
        rt, emb, gate, hidx, htot = wavlm
    check(_L().dph_softmax_bwd(ptr(P), ptr(dPd), lds, ptr(dS), ldp, B, H, T, drop_p, seed, ptr(rt), ptr(emb), ptr(gate),
                               ptr(hidx), htot, ptr(dgate), ptr(demb), num_buckets, dtype_code(P), stream_ptr()),
          "dph_softmax_bwd")


def wavlm_gate_fwd(x, Wg, bg, cst, gate, ga, gb, B, T, Htot, Dh):
    check(_L().dph_wavlm_gate_fwd(ptr(x), ptr(Wg), ptr(bg), ptr(cst), ptr(gate), ptr(ga), ptr(gb), B, T, Htot, Dh,
                                  dtype_code(x), stream_ptr()), "dph_wavlm_gate_fwd")


def wavlm_gate_bwd(x, Wg, cst, ga, gb, dgate, dx, dWg, dbg, dcst, B, T, Htot, Dh):
    check(_L().dph_wavlm_gate_bwd(ptr(x), ptr(Wg), ptr(cst), ptr(ga), ptr(gb), ptr(dgate), ptr(dx), ptr(dWg), ptr(dbg),
                                  ptr(dcst), B, T, Htot, Dh, dtype_code(x), stream_ptr()), "dph_wavlm_gate_bwd")


def attn_fwd(qkv, ctx, ctx_un, lse, B, H, T, scale, lengths=None, drop_p=0.0, seed=0, wavlm=None, head_mask=None):
    """Fused attention forward (head_dim 64, bf16): qkv (B*T, 3*H*64) laid out [k | v | q] -> ctx (B*T, H*64)."""
    rt = emb = gate = hidx = None
    htot = H
    if wavlm is not None:
        rt, emb, gate, hidx, htot = wavlm
    check(_L().dph_attn_fwd(ptr(qkv), ptr(ctx), ptr(ctx_un), ptr(lse), B, H, T, scale, ptr(lengths), ptr(rt), ptr(emb),
                            ptr(gate), ptr(hidx), htot, drop_p, seed, ptr(head_mask), stream_ptr()), "dph_attn_fwd")


def attn_bwd(qkv, dctx, o_un, lse, delta, dq32, dqkv, dmask_h, B, H, T, scale, lengths=None, drop_p=0.0, seed=0,
             wavlm=None, head_mask=None, dgate=None, demb=None, num_buckets=0):
    rt = emb = gate = hidx = None
    htot = H
    if wavlm is not None:
        rt, emb, gate, hidx, htot = wavlm
    check(_L().dph_attn_bwd(ptr(qkv), ptr(dctx), ptr(o_un), ptr(lse), ptr(delta), ptr(dq32), ptr(dqkv), ptr(dmask_h), B, H,
                            T, scale, ptr(lengths), ptr(rt), ptr(emb), ptr(gate), ptr(hidx), htot, drop_p, seed,
                            ptr(head_mask), ptr(dgate), ptr(demb), num_buckets, stream_ptr()), "dph_attn_bwd")


# ----------------------------------------------------------------------------------------------------------------
# distillation loss
# ----------------------------------------------------------------------------------------------------------------
def distill_loss_fwd(a, b, rowstat, acc, out, rows, D, w2, w1, wc, cos_type, first, last, total_rows):
    check(_L().dph_distill_loss_fwd(ptr(a), ptr(b), ptr(rowstat), ptr(acc), ptr(out), rows, D, w2, w1, wc, cos_type,
                                    int(first), int(last), total_rows, dtype_code(a), stream_ptr()),
          "dph_distill_loss_fwd")


def distill_loss_bwd(a, b, rowstat, upstream, da, rows, total_rows, D, w2, w1, wc, cos_type):
    check(_L().dph_distill_loss_bwd(ptr(a), ptr(b), ptr(rowstat), ptr(upstream), ptr(da), rows, total_rows, D, w2, w1,
                                    wc, cos_type, dtype_code(a), stream_ptr()), "dph_distill_loss_bwd")


# ----------------------------------------------------------------------------------------------------------------
# weight preparation
# ----------------------------------------------------------------------------------------------------------------
def conv_weight_prep(w, wf, wd, Cout, Cin, k, perm):
    arr = (ctypes.c_int * 16)(*([int(x) for x in perm] + [0] * (16 - len(perm))))
    out = wf if wf is not None else wd
    check(_L().dph_conv_weight_prep(ptr(w), ptr(wf), ptr(wd), Cout, Cin, k, arr, dtype_code(out), stream_ptr()),
          "dph_conv_weight_prep")


def conv_wgrad_unpermute(src, dw, Cout, Cin, k):
    check(_L().dph_conv_wgrad_unpermute(ptr(src), ptr(dw), Cout, Cin, k, stream_ptr()), "dph_conv_wgrad_unpermute")


def posconv_weight_prep(v, g, norm, wn, wp, wdp, D, G, Cg, Cgp, K):
    out = wp if wp is not None else wdp
    code = dtype_code(out) if out is not None else 1
    check(_L().dph_posconv_weight_prep(ptr(v), ptr(g), ptr(norm), ptr(wn), ptr(wp), ptr(wdp), D, G, Cg, Cgp, K, code,
                                       stream_ptr()), "dph_posconv_weight_prep")


def posconv_weight_bwd(dwp, v, g, norm, scratch, dv, dg, D, G, Cg, Cop, Cgp, K):
    check(_L().dph_posconv_weight_bwd(ptr(dwp), ptr(v), ptr(g), ptr(norm), ptr(scratch), ptr(dv), ptr(dg), D, G, Cg, Cop,
                                      Cgp, K, stream_ptr()), "dph_posconv_weight_bwd")


# ----------------------------------------------------------------------------------------------------------------
# optimizer pieces
# ----------------------------------------------------------------------------------------------------------------
def sumsq(x, out):
    check(_L().dph_sumsq(ptr(x), x.numel(), ptr(out), stream_ptr()), "dph_sumsq")


def counter_inc(c):
    check(_L().dph_counter_inc(ptr(c), stream_ptr()), "dph_counter_inc")


def adamw(p, g, m, v, lr_dev, beta1, beta2, eps, wd, step_dev, gnorm_sq, max_norm, p_bf16, zero_grad):
    check(_L().dph_adamw(ptr(p), ptr(g), ptr(m), ptr(v), p.numel(), ptr(lr_dev), beta1, beta2, eps, wd, ptr(step_dev),
                         ptr(gnorm_sq), max_norm, ptr(p_bf16), int(zero_grad), stream_ptr()), "dph_adamw")


def scale_(x, a):
    check(_L().dph_scale(ptr(x), x.numel(), a, stream_ptr()), "dph_scale")


def multimem_allreduce(mc_addr, n, rank, world, scale, ctas=0):
    """`mc_addr`: integer multicast address of the slice (symmetric memory handle's multicast_ptr + byte offset)."""
    check(_L().dph_multimem_allreduce(ctypes.c_void_p(int(mc_addr)), int(n), int(rank), int(world), float(scale), int(ctas),
                                      stream_ptr()), "dph_multimem_allreduce")


# ----------------------------------------------------------------------------------------------------------------
# HardConcrete
# ----------------------------------------------------------------------------------------------------------------
def hardconcrete_fwd(la, u_in, u_out, mask, n, beta, l, r, eps, seed):
    check(_L().dph_hardconcrete_fwd(ptr(la), ptr(u_in), ptr(u_out), ptr(mask), n, beta, l, r, eps, seed, stream_ptr()),
          "dph_hardconcrete_fwd")


def hardconcrete_bwd(la, u, dmask, dla, n, beta, l, r):
    check(_L().dph_hardconcrete_bwd(ptr(la), ptr(u), ptr(dmask), ptr(dla), n, beta, l, r, stream_ptr()),
          "dph_hardconcrete_bwd")


def l0_norms(la_base, seg_off, seg_len, nseg, bias, norms):
    check(_L().dph_l0_norms(ptr(la_base), ptr(seg_off), ptr(seg_len), nseg, bias, ptr(norms), stream_ptr()),
          "dph_l0_norms")


def num_params(prog, norms, nseg, out_n, dnorm):
    check(_L().dph_num_params(ptr(prog), ptr(norms), nseg, ptr(out_n), ptr(dnorm), stream_ptr()), "dph_num_params")


def lagrangian(n_hat, lambdas, n0, target, out):
    check(_L().dph_lagrangian(ptr(n_hat), ptr(lambdas), n0, ptr(target), ptr(out), stream_ptr()), "dph_lagrangian")


def l0_norms_bwd(la_base, dla_base, seg_off, seg_len, nseg, bias, dnorm, upstream, coef):
    check(_L().dph_l0_norms_bwd(ptr(la_base), ptr(dla_base), ptr(seg_off), ptr(seg_len), nseg, bias, ptr(dnorm),
                                ptr(upstream), ptr(coef), stream_ptr()), "dph_l0_norms_bwd")


_seed_counter = [0x5DEECE66D]


def next_seed():
    """Host-side stream of kernel RNG seeds, re-seeded from torch's generator so torch.manual_seed controls it."""
    _seed_counter[0] = (_seed_counter[0] * 6364136223846793005 + 1442695040888963407) & 0xFFFFFFFFFFFFFFFF
    return _seed_counter[0] >> 1


def rng_set(value):
    check(_L().dph_rng_set(int(value) & 0xFFFFFFFFFFFFFFFF, stream_ptr()), "dph_rng_set")


def rng_advance():
    check(_L().dph_rng_advance(stream_ptr()), "dph_rng_advance")


def rng_get(device):
    """Current device-side RNG epoch as a Python int (synchronises; checkpointing only)."""
    out = torch.zeros(1, dtype=torch.int64, device=device)
    check(_L().dph_rng_get(ptr(out), stream_ptr()), "dph_rng_get")
    return int(out.cpu()[0])


def seed_state():
    return int(_seed_counter[0])


def set_seed_state(v):
    _seed_counter[0] = int(v) & 0xFFFFFFFFFFFFFFFF


def manual_seed(seed):
    _seed_counter[0] = (int(seed) * 0x9E3779B97F4A7C15 + 0x5DEECE66D) & 0xFFFFFFFFFFFFFFFF


class _HardConcreteSample(torch.autograd.Function):
    """mask = clamp(sigmoid((logit(u) + log_alpha) / beta) * (r - l) + l, 0, 1)   (hardconcrete.py:96-99)"""

    @staticmethod
    def forward(ctx, log_alpha, u, beta, l, r, eps):
        la = log_alpha.detach().contiguous()
        n = la.numel()
        mask = torch.empty_like(la)
        if u is None:
            u_used = torch.empty_like(la)
            hardconcrete_fwd(la, None, u_used, mask, n, beta, l, r, eps, next_seed())
        else:
            u_used = u.detach().to(la.device, torch.float32).contiguous()
            hardconcrete_fwd(la, u_used, None, mask, n, beta, l, r, eps, 0)
        ctx.save_for_backward(la, u_used)
        ctx.hp = (beta, l, r)
        return mask

    @staticmethod
    def backward(ctx, dmask):
        la, u = ctx.saved_tensors
        beta, l, r = ctx.hp
        dla = zeros(la.shape, torch.float32, la.device)
        hardconcrete_bwd(la, u, dmask.contiguous().float(), dla, la.numel(), beta, l, r)
        return dla, None, None, None, None, None


def hardconcrete_sample(log_alpha, u, beta, limit_l, limit_r, eps=1e-6):
    return _HardConcreteSample.apply(log_alpha, u, float(beta), float(limit_l), float(limit_r), float(eps))


class _L0Norm(torch.autograd.Function):
    """sum(sigmoid(log_alpha + bias))   (hardconcrete.py:76-83)"""

    @staticmethod
    def forward(ctx, log_alpha, bias):
        la = log_alpha.detach().contiguous()
        dev = la.device
        seg_off = torch.zeros(1, dtype=torch.int64, device=dev)
        seg_len = torch.full((1,), la.numel(), dtype=torch.int32, device=dev)
        out = torch.empty(1, dtype=torch.float32, device=dev)
        l0_norms(la, seg_off, seg_len, 1, bias, out)
        ctx.save_for_backward(la, seg_off, seg_len)
        ctx.bias = bias
        return out[0]

    @staticmethod
    def backward(ctx, g):
        la, seg_off, seg_len = ctx.saved_tensors
        dla = zeros(la.shape, torch.float32, la.device)
        one = torch.ones(1, dtype=torch.float32, device=la.device)
        l0_norms_bwd(la, dla, seg_off, seg_len, 1, ctx.bias, one, g.reshape(1).float().contiguous(), None)
        return dla, None


def l0_norm(log_alpha, bias):
    return _L0Norm.apply(log_alpha, float(bias))
