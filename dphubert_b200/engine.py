"""Native execution engine behind the `wav2vec2` module mirror.

Per model a `Plan` owns
  * one flat fp32 *parameter arena* (all nn.Parameters are re-pointed to views of it; GEMM weights first, then
    vectors, then every HardConcrete `log_alpha` contiguously) and a matching *gradient arena* — so that the
    optimizer, the gradient all-reduce, the bf16 weight refresh and the HardConcrete kernels are single launches
    over contiguous memory;
  * the derived GEMM operands (bf16 weight copy, tap-major conv weights, weight-normalised positional conv);
  * the static tables of the expected-size formula.
The forward/backward of the three blocks (conv feature extractor, encoder, distillation heads + loss) are plain
sequences of C-ABI launches wrapped in one autograd node each.  Parameter gradients are written straight into the
gradient arena (`p.grad` is a view of it); the autograd nodes return None for parameters.
"""
import math
from typing import List, Optional

import torch

from . import _lib, ops
from . import gemm as G
from .gemm import View

ALIGN = 8  # elements; keeps every bf16 view 16-byte aligned


def _align(n):
    return (n + ALIGN - 1) // ALIGN * ALIGN


class DphShapeError(_lib.DphError):
    pass


def _require_cuda(t, what):
    if not t.is_cuda:
        raise _lib.DphError(
            "dphubert_b200: %s is on %s — the model runs only on CUDA through the native kernels "
            "(there is no CPU fallback); move the model and inputs to a GPU" % (what, t.device)
        )


# ----------------------------------------------------------------------------------------------------------------
# Plan
# ----------------------------------------------------------------------------------------------------------------
class Plan:
    def __init__(self, model):
        self.model = model
        params = list(model.named_parameters())
        if not params:
            raise _lib.DphError("model has no parameters")
        self.device = params[0][1].device
        _require_cuda(params[0][1], "the model")
        fe = model.feature_extractor
        enc = model.encoder
        tr = enc.transformer

        lin, la = [], []
        for layer in tr.layers:
            att, ff = layer.attention, layer.feed_forward
            if att is not None:
                lin += [att.k_proj.weight, att.v_proj.weight, att.q_proj.weight, att.out_proj.weight]
            if ff is not None:
                lin += [ff.intermediate_dense.weight, ff.output_dense.weight]
        lin.append(enc.feature_projection.projection.weight)
        if model.aux is not None:
            lin.append(model.aux.weight)
        for blk in fe.conv_layers:
            if blk.hard_concrete is not None:
                la.append(blk.hard_concrete.log_alpha)
        for layer in tr.layers:
            att, ff = layer.attention, layer.feed_forward
            if att is not None:
                for hc in (att.hard_concrete_for_heads, att.hard_concrete_for_layer):
                    if hc is not None:
                        la.append(hc.log_alpha)
            if ff is not None:
                for hc in (ff.hard_concrete_for_intermediate, ff.hard_concrete_for_layer):
                    if hc is not None:
                        la.append(hc.log_alpha)
        special = {id(p) for p in lin} | {id(p) for p in la}
        rest = [p for _, p in params if id(p) not in special]
        ordered = lin + rest + la
        assert len({id(p) for p in ordered}) == len(params)

        self.offsets = {}
        off = 0
        for p in lin:
            self.offsets[id(p)] = off
            off += _align(p.numel())
        self.lin_total = off
        for p in rest:
            self.offsets[id(p)] = off
            off += _align(p.numel())
        self.la_off = off
        for p in la:
            self.offsets[id(p)] = off
            off += _align(p.numel())
        self.total = off
        self.la_n = off - self.la_off
        self.la_params = la

        dev = self.device
        self.arena = ops.zeros((self.total,), torch.float32, dev)
        self.any_grad = any(p.requires_grad for _, p in params)
        self.garena = ops.zeros((self.total,), torch.float32, dev) if self.any_grad else None
        with torch.no_grad():
            for p in ordered:
                o, n = self.offsets[id(p)], p.numel()
                view = self.arena[o:o + n].view(p.shape)
                view.copy_(p.data)  # one-time setup copy
                p.data = view
        self.params = ordered
        self.bind_grads()

        self.dtype = None
        self.lin16 = None
        self.conv_w = {}     # layer index -> (wf, wd) in compute dtype (bf16 mode only)
        self.pos_w = None    # (norm, wp, wdp)
        self.dirty = True
        self.version_key = None
        self._hc_tables = None
        self._rel_tables = {}

    # ------------------------------------------------------------------ parameter / gradient views
    def view(self, p):
        o = self.offsets[id(p)]
        return self.arena[o:o + p.numel()].view(p.shape)

    def gview(self, p):
        o = self.offsets[id(p)]
        return self.garena[o:o + p.numel()].view(p.shape)

    def bind_grads(self):
        """(Re-)attach `p.grad` to the gradient arena; a grad that was set to None counts as zero."""
        if self.garena is None:
            return
        for p in self.params:
            if not p.requires_grad:
                continue
            g = self.gview(p)
            if p.grad is None:
                p.grad = g
            elif p.grad.data_ptr() != g.data_ptr():
                g.copy_(p.grad)
                p.grad = g

    def grads_ok(self):
        if self.garena is None:
            return True
        base, end = self.garena.data_ptr(), self.garena.data_ptr() + self.total * 4
        for p in self.params:
            if p.requires_grad and (p.grad is None or not (base <= p.grad.data_ptr() < end)):
                return False
        return True

    def prepare_backward(self):
        """Called at the top of every native backward: make sure p.grad still points into the arena. If a foreign
        optimizer did zero_grad(set_to_none=True) the arena slice is zeroed (that is what None means)."""
        if self.garena is None or self.grads_ok():
            return
        for p in self.params:
            if not p.requires_grad:
                continue
            g = self.gview(p)
            if p.grad is None:
                ops.memset(g)
                p.grad = g
            elif p.grad.data_ptr() != g.data_ptr():
                g.copy_(p.grad)
                p.grad = g

    def still_bound(self):
        base, end = self.arena.data_ptr(), self.arena.data_ptr() + self.total * 4
        return all(base <= p.data_ptr() < end for p in self.params)

    # ------------------------------------------------------------------ derived weights
    def _vkey(self):
        return sum(p._version for p in self.params)

    def mark_dirty(self):
        self.dirty = True

    def refresh(self, dtype):
        """Rebuild the derived GEMM operands if any parameter changed (or the compute dtype did)."""
        key = self._vkey()
        if not self.dirty and self.dtype == dtype and key == self.version_key:
            return
        model = self.model
        dev = self.device
        bf = dtype == torch.bfloat16
        if bf:
            if self.lin16 is None:
                self.lin16 = torch.empty(self.lin_total, dtype=torch.bfloat16, device=dev)
            ops.cast(self.arena[: self.lin_total], self.lin16)
            for i, blk in enumerate(model.feature_extractor.conv_layers):
                if i == 0:
                    continue
                w = blk.conv.weight
                Cout, Cin, k = w.shape
                if Cin % 8 or Cout % 8:
                    raise DphShapeError(
                        "bf16 tensor-core mode needs conv channel counts that are multiples of 8 (layer %d has "
                        "%d -> %d); use compute dtype float32 for ragged pruned models" % (i, Cin, Cout))
                if i not in self.conv_w or self.conv_w[i][0].dtype != dtype or self.conv_w[i][0].numel() != w.numel():
                    self.conv_w[i] = (torch.empty(Cout, k, Cin, dtype=dtype, device=dev),
                                      torch.empty(k, Cin, Cout, dtype=dtype, device=dev))
                perm = _dgrad_tap_perm(k, blk.stride)
                ops.conv_weight_prep(self.view(w), self.conv_w[i][0], self.conv_w[i][1], Cout, Cin, k, perm)
        pc = model.encoder.transformer.pos_conv_embed
        v, g = pc.conv.weight_v, pc.conv.weight_g
        D, Cg, K = v.shape
        Gn = D // Cg
        Cgp = _align_to(Cg, 64) if bf else Cg
        if self.pos_w is None or self.pos_w[1].dtype != dtype:
            norm = torch.empty(K, dtype=torch.float32, device=dev)
            wp = ops.zeros((Gn, Cg, K, Cgp), dtype, dev)
            wdp = ops.zeros((Gn, Cg, K, Cgp), dtype, dev)
            self.pos_w = (norm, wp, wdp, Cgp)
        norm, wp, wdp, _ = self.pos_w
        ops.posconv_weight_prep(self.view(v), self.view(g), norm, None, wp, wdp, D, Gn, Cg, Cgp, K)
        self.dtype = dtype
        self.dirty = False
        self.version_key = key

    def w16(self, p):
        """bf16 copy of a Linear weight (view into the bf16 arena)."""
        o = self.offsets[id(p)]
        return self.lin16[o:o + p.numel()].view(p.shape)


def _align_to(n, a):
    return (n + a - 1) // a * a


def _dgrad_tap_perm(k, stride):
    """Order of taps in the dgrad weight copy: taps of the same output phase (tap % stride) are made contiguous,
    within a phase in decreasing tap order (row t'-j of the upstream gradient pairs with tap phase + j*stride)."""
    order = []
    for phase in range(stride):
        taps = [t for t in range(k) if t % stride == phase]
        order += sorted(taps, reverse=True)
    perm = [0] * k
    for pos, tap in enumerate(order):
        perm[tap] = pos
    return perm


_PLANS = {}


def get_plan(model) -> Plan:
    plan = _PLANS.get(id(model))
    if plan is not None and plan.model is model and plan.still_bound() and len(plan.params) == sum(
            1 for _ in model.parameters()):
        return plan
    plan = Plan(model)
    _PLANS[id(model)] = plan
    return plan


def invalidate(model):
    _PLANS.pop(id(model), None)


# ----------------------------------------------------------------------------------------------------------------
# HardConcrete plumbing: all gates of a model are sampled / reduced in single launches over the log_alpha region
# ----------------------------------------------------------------------------------------------------------------
class HcTables:
    """Static description of a model's gates: segments inside the log_alpha region and the size-formula program
    consumed by `dph_num_params` (layout documented in csrc/hardconcrete.cu)."""

    def __init__(self, plan: Plan):
        model = plan.model
        dev = plan.device
        segs = []  # (param, offset relative to la region)
        seg_index = {}

        def seg_of(hc):
            if hc is None:
                return -1
            p = hc.log_alpha
            if id(p) not in seg_index:
                seg_index[id(p)] = len(segs)
                segs.append((p, plan.offsets[id(p)] - plan.la_off))
            return seg_index[id(p)]

        fe = model.feature_extractor
        enc = model.encoder
        tr = enc.transformer
        prog = [len(fe.conv_layers)]
        hcs = []
        for blk in fe.conv_layers:
            prog += [seg_of(blk.hard_concrete), blk.kernel_size, int(blk.conv.bias is not None),
                     int(blk.layer_norm is not None), blk.conv.out_channels]
            if blk.hard_concrete is not None:
                hcs.append(blk.hard_concrete)
        D = tr.pos_conv_embed.embed_dim
        static_extra = sum(p.numel() for p in tr.pos_conv_embed.parameters()) + 2 * D
        prog += [enc.feature_projection.projection.out_features, static_extra, D, len(tr.layers)]
        for layer in tr.layers:
            att, ff = layer.attention, layer.feed_forward
            if att is not None:
                prog += [1, seg_of(att.hard_concrete_for_heads), att.num_heads, att.head_dim,
                         seg_of(att.hard_concrete_for_layer)]
                hcs += [h for h in (att.hard_concrete_for_heads, att.hard_concrete_for_layer) if h is not None]
            else:
                prog += [0, -1, 0, 0, -1]
            if ff is not None:
                prog += [1, seg_of(ff.hard_concrete_for_intermediate), ff.intermediate_dense.out_features,
                         seg_of(ff.hard_concrete_for_layer)]
                hcs += [h for h in (ff.hard_concrete_for_intermediate, ff.hard_concrete_for_layer) if h is not None]
            else:
                prog += [0, -1, 0, -1]
        self.hcs = hcs
        self.nseg = len(segs)
        self.segs = segs
        if self.nseg:
            hc0 = hcs[0]
            self.bias, self.beta, self.l, self.r, self.eps = hc0.bias, hc0.beta, hc0.limit_l, hc0.limit_r, hc0.eps
            for h in hcs:
                assert (h.bias, h.beta, h.limit_l, h.limit_r) == (self.bias, self.beta, self.l, self.r), \
                    "all HardConcrete gates of a model must share temperature / stretch"
            self.seg_off = torch.tensor([o for _, o in segs], dtype=torch.int64).to(dev)
            self.seg_len = torch.tensor([p.numel() for p, _ in segs], dtype=torch.int32).to(dev)
        self.prog = torch.tensor(prog, dtype=torch.int32).to(dev)


def hc_tables(plan: Plan) -> HcTables:
    if plan._hc_tables is None:
        plan._hc_tables = HcTables(plan)
    return plan._hc_tables


class _SampleMasks(torch.autograd.Function):
    """All HardConcrete masks of a model in one launch (hardconcrete.py:92-99 x 31 modules)."""

    @staticmethod
    def forward(ctx, plan, u_flat, *log_alphas):
        tb = hc_tables(plan)
        la = plan.arena[plan.la_off:]
        masks = torch.empty(plan.la_n, dtype=torch.float32, device=plan.device)
        if u_flat is None:
            u_used = torch.empty(plan.la_n, dtype=torch.float32, device=plan.device)
            ops.hardconcrete_fwd(la, None, u_used, masks, plan.la_n, tb.beta, tb.l, tb.r, tb.eps, ops.next_seed())
        else:
            u_used = u_flat
            ops.hardconcrete_fwd(la, u_used, None, masks, plan.la_n, tb.beta, tb.l, tb.r, tb.eps, 0)
        ctx.plan = plan
        ctx.u = u_used
        return masks

    @staticmethod
    def backward(ctx, dmasks):
        plan = ctx.plan
        tb = hc_tables(plan)
        plan.prepare_backward()
        la = plan.arena[plan.la_off:]
        dla = plan.garena[plan.la_off:]
        ops.hardconcrete_bwd(la, ctx.u, dmasks.contiguous(), dla, plan.la_n, tb.beta, tb.l, tb.r)
        return (None, None) + (None,) * len(plan.la_params)


def sample_masks(plan: Plan):
    """Returns {id(hc module): mask tensor view} for the current forward (training: sampled; eval: compiled)."""
    tb = hc_tables(plan)
    if tb.nseg == 0:
        return {}
    model = plan.model
    out = {}
    if model.training:
        u_flat = None
        if any(h.noise is not None for h in tb.hcs):
            # parity hook: the caller injected the uniform noise of (some) gates
            u_flat = torch.empty(plan.la_n, dtype=torch.float32, device=plan.device).uniform_(tb.eps, 1 - tb.eps)
            for h in tb.hcs:
                if h.noise is not None:
                    o = plan.offsets[id(h.log_alpha)] - plan.la_off
                    u_flat[o:o + h.n_in].copy_(h.noise.to(plan.device, torch.float32))
                    h.noise = None
        for h in tb.hcs:
            h.compiled_mask = None
        masks = _SampleMasks.apply(plan, u_flat, *plan.la_params)
        for h in tb.hcs:
            o = plan.offsets[id(h.log_alpha)] - plan.la_off
            out[id(h)] = masks[o:o + h.n_in]
    else:
        for h in tb.hcs:
            out[id(h)] = h().detach().to(plan.device, torch.float32).contiguous()
    return out


class _NumParams(torch.autograd.Function):
    """Differentiable expected parameter count (model.py:109-113 and the get_num_params tree)."""

    @staticmethod
    def forward(ctx, plan, *log_alphas):
        tb = hc_tables(plan)
        dev = plan.device
        la = plan.arena[plan.la_off:]
        norms = torch.empty(tb.nseg, dtype=torch.float32, device=dev)
        out = torch.empty(1, dtype=torch.float32, device=dev)
        dnorm = torch.empty(tb.nseg, dtype=torch.float32, device=dev)
        ops.l0_norms(la, tb.seg_off, tb.seg_len, tb.nseg, tb.bias, norms)
        ops.num_params(tb.prog, norms, tb.nseg, out, dnorm)
        ctx.plan = plan
        ctx.dnorm = dnorm
        return out[0]

    @staticmethod
    def backward(ctx, g):
        plan = ctx.plan
        tb = hc_tables(plan)
        plan.prepare_backward()
        la = plan.arena[plan.la_off:]
        dla = plan.garena[plan.la_off:]
        ops.l0_norms_bwd(la, dla, tb.seg_off, tb.seg_len, tb.nseg, tb.bias, ctx.dnorm,
                         g.reshape(1).float().contiguous(), None)
        return (None,) + (None,) * len(plan.la_params)


def fused_num_params(model):
    """Expected size through the fused kernels when the model has gates and lives on a GPU; None otherwise (the
    caller then falls back to the exact integer bookkeeping, which involves no tensor math)."""
    p0 = next(model.parameters())
    if not p0.is_cuda:
        return None
    plan = get_plan(model)
    tb = hc_tables(plan)
    if tb.nseg == 0:
        return None
    return _NumParams.apply(plan, *plan.la_params)


# ----------------------------------------------------------------------------------------------------------------
# Convolutional feature extractor (components.py:94-120, 158-185)
# ----------------------------------------------------------------------------------------------------------------
def _frames(n, k, s):
    return (n - k) // s + 1 if n >= k else 0


def _conv_weight_views(plan, blk, i, dtype):
    """(forward B operand, list of per-phase dgrad B operands) for conv layer i >= 1."""
    w = blk.conv.weight
    Cout, Cin, k = w.shape
    s = blk.stride
    phases = []
    if dtype == torch.bfloat16:
        wf, wd = plan.conv_w[i]
        fwd = View(wf, Cout, k * Cin, k * Cin)
        pos = 0
        for ph in range(s):
            J = len([t for t in range(k) if t % s == ph])
            if J:
                phases.append((ph, J, View(wd, Cin, J * Cout, Cout, kin=Cout, kxs=Cin * Cout, offset=pos * Cin * Cout)))
            pos += J
    else:
        wv = plan.view(w)
        fwd = View(wv, Cout, k * Cin, Cin * k, cs=k, kin=Cin, kxs=1)
        for ph in range(s):
            J = len([t for t in range(k) if t % s == ph])
            if J:
                phases.append((ph, J, View(wv, Cin, J * Cout, k, cs=Cin * k, kin=Cout, kxs=-s, offset=ph + (J - 1) * s)))
    return fwd, phases


def conv_stack_fwd(plan: Plan, wave, masks, save: bool):
    """wave (B, T0) fp32 -> features (B, T, C) in the compute dtype. Returns (features, saved-state dict)."""
    model = plan.model
    fe = model.feature_extractor
    dt = model.compute_dtype
    dev = plan.device
    B, T0 = wave.shape
    wave = wave.contiguous()
    if wave.dtype != torch.float32:
        raise _lib.DphError("waveforms must be float32")
    st = {"wave": wave, "layers": [], "B": B}
    x = None
    Tin, Cin = T0, 1
    n_layers = len(fe.conv_layers)
    for i, blk in enumerate(fe.conv_layers):
        conv = blk.conv
        Cout, k, s = conv.out_channels, blk.kernel_size, blk.stride
        Tout = _frames(Tin, k, s)
        if Tout <= 0:
            raise ValueError("input too short for the convolutional feature extractor")
        mask = masks.get(id(blk.hard_concrete)) if blk.hard_concrete is not None else None
        colmask = mask
        if i == n_layers - 1:
            # x * dummy_weight after the last block (components.py:183-184) folded into the channel mask
            dummy = plan.view(fe.dummy_weight)
            if mask is None:
                colmask = dummy
            else:
                colmask = torch.empty_like(dummy)
                ops.vec_mul(mask, dummy, colmask)
        ln = blk.layer_norm
        is_gn = isinstance(ln, torch.nn.GroupNorm)
        bias = plan.view(conv.bias) if conv.bias is not None else None
        rec = {"blk": blk, "i": i, "Tin": Tin, "Tout": Tout, "Cin": Cin, "Cout": Cout, "colmask": colmask,
               "has_mask": mask is not None, "x_in": x}
        y = torch.empty(B, Tout, Cout, dtype=dt, device=dev)
        if i == 0:
            if Cin != 1:
                raise DphShapeError("first conv layer must have one input channel")
            w = plan.view(conv.weight)
            if is_gn:
                if ln.num_groups != Cout:
                    raise DphShapeError("GroupNorm must have one group per channel")
                moments = torch.empty(B * 65, dtype=torch.float64, device=dev)
                mean = torch.empty(B * Cout, dtype=torch.float32, device=dev)
                rstd = torch.empty(B * Cout, dtype=torch.float32, device=dev)
                ops.conv0_stats(wave, w, bias, moments, mean, rstd, B, Cout, k, s, Tout, ln.eps)
                ops.conv0_fwd(wave, w, bias, plan.view(ln.weight), plan.view(ln.bias), mean, rstd, colmask, y, B, Cout,
                              k, s, Tout, 1, 1)
                rec.update(mean=mean, rstd=rstd, mode="gn")
            elif ln is not None:
                z = torch.empty(B, Tout, Cout, dtype=dt, device=dev)
                ops.conv0_fwd(wave, w, bias, None, None, None, None, None, z, B, Cout, k, s, Tout, 0, 0)
                mean = torch.empty(B * Tout, dtype=torch.float32, device=dev)
                rstd = torch.empty(B * Tout, dtype=torch.float32, device=dev)
                ops.layernorm_fwd(z, plan.view(ln.weight), plan.view(ln.bias), y, B * Tout, Cout, ln.eps, mean, rstd,
                                  colmask=colmask, act=1)
                rec.update(z=z, mean=mean, rstd=rstd, mode="ln")
            else:
                ops.conv0_fwd(wave, w, bias, None, None, None, None, colmask, y, B, Cout, k, s, Tout, 0, 1)
                rec.update(mode="none")
        else:
            fwd_w, _ = _conv_weight_views(plan, blk, i, dt)
            a = View(x, Tout, k * Cin, s * Cin, sb0=Tin * Cin)
            if ln is None:
                z = torch.empty(B, Tout, Cout, dtype=dt, device=dev) if save else None
                G.gemm(a, fwd_w, y, Tout, Cout, k * Cin, nb0=B, c_sb0=Tout * Cout, bias=bias, act=1, colmask=colmask,
                       c2=z)
                rec.update(z=z, mode="none")
            else:
                if is_gn:
                    raise DphShapeError("GroupNorm is only supported on the first conv layer (as in the reference)")
                z = torch.empty(B, Tout, Cout, dtype=dt, device=dev)
                G.gemm(a, fwd_w, z, Tout, Cout, k * Cin, nb0=B, c_sb0=Tout * Cout, bias=bias)
                mean = torch.empty(B * Tout, dtype=torch.float32, device=dev)
                rstd = torch.empty(B * Tout, dtype=torch.float32, device=dev)
                ops.layernorm_fwd(z, plan.view(ln.weight), plan.view(ln.bias), y, B * Tout, Cout, ln.eps, mean, rstd,
                                  colmask=colmask, act=1)
                rec.update(z=z, mean=mean, rstd=rstd, mode="ln")
        rec["y"] = y
        st["layers"].append(rec)
        x = y
        Tin, Cin = Tout, Cout
    return x, st
