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
GRAD_TAIL = 8  # spare floats behind the gradient arena: the training step keeps d loss / d lambda there


def _align(n):
    return (n + ALIGN - 1) // ALIGN * ALIGN


class DphShapeError(_lib.DphError):
    pass


def _is_native(t):
    return t.is_cuda


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

        # ---- arena order = order in which the backward pass COMPLETES gradients, so that the data-parallel gradient
        # mean can be issued as a few contiguous slices while the rest of the backward still runs (trainer.py):
        #   [layer 0 | layer 1 | ... | layer L-1 | projection, aux | transformer-level / positional conv / projection
        #    vectors | conv n-1 .. conv 1 | conv 0, dummy_weight | every log_alpha]  (+ GRAD_TAIL floats in the gradient
        #   arena only, for the caller's Lagrange multipliers).
        # Inside a layer the Linear weights come first (K, V, Q back to back: fused projection), then its vectors.
        def gates_of(mod):
            return [m.log_alpha for m in mod.modules() if hasattr(m, "log_alpha") and hasattr(m, "n_in")]

        la = []
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
        la_ids = {id(p) for p in la}
        groups = []   # per transformer layer
        for layer in tr.layers:
            att, ff = layer.attention, layer.feed_forward
            ws = []
            if att is not None:
                ws += [att.k_proj.weight, att.v_proj.weight, att.q_proj.weight, att.out_proj.weight]
            if ff is not None:
                ws += [ff.intermediate_dense.weight, ff.output_dense.weight]
            wid = {id(p) for p in ws}
            groups.append(ws + [p for p in layer.parameters() if id(p) not in wid and id(p) not in la_ids])
        lin_tail = [enc.feature_projection.projection.weight]
        if model.aux is not None:
            lin_tail.append(model.aux.weight)
        conv_hi, conv_lo = [], []
        blocks = list(fe.conv_layers)
        for i in range(len(blocks) - 1, 0, -1):
            conv_hi += [p for p in blocks[i].parameters() if id(p) not in la_ids]
        if blocks:
            conv_lo += [p for p in blocks[0].parameters() if id(p) not in la_ids]
        conv_lo += [p for p in fe.parameters() if id(p) not in la_ids and
                    all(p is not q for q in conv_hi + conv_lo)]   # dummy_weight
        placed = {id(p) for g_ in groups for p in g_} | {id(p) for p in lin_tail + conv_hi + conv_lo} | la_ids
        mid = [p for _, p in params if id(p) not in placed]
        ordered = [p for g_ in groups for p in g_] + lin_tail + mid + conv_hi + conv_lo + la
        assert len({id(p) for p in ordered}) == len(params) == len(ordered)

        self.offsets = {}
        off = 0
        marks = {}

        def place(ps):
            nonlocal off
            start = off
            for p in ps:
                self.offsets[id(p)] = off
                off += _align(p.numel())
            return (start, off) if off > start else None

        self.layer_lin_range = [place(g_) for g_ in groups]
        enc_start = off
        place(lin_tail)
        self.lin_total = off      # the bf16 mirror covers [0, lin_total): every GEMM weight (and the layers' vectors)
        place(mid)
        marks["encoder"] = (enc_start, off)
        marks["conv_hi"] = place(conv_hi)
        lo_start = off
        place(conv_lo)
        self.la_off = off
        place(la)
        self.total = off
        self.la_n = off - self.la_off
        self.la_params = la
        marks["tail"] = (lo_start, self.total + GRAD_TAIL)
        # gradient slices by completion event: int (transformer layer), "encoder" (everything between the transformer
        # layers and the conv stack), "conv_hi" (conv layers n-1 .. 1), "tail" (conv 0, gates, the caller's tail)
        self.grad_ranges = dict(marks)
        for li, r in enumerate(self.layer_lin_range):
            self.grad_ranges[li] = r

        dev = self.device
        self._vcache, self._gcache, self._wcache = {}, {}, {}
        self.arena = ops.zeros((self.total,), torch.float32, dev)
        self.any_grad = any(p.requires_grad for _, p in params)
        self.garena_full = ops.zeros((self.total + GRAD_TAIL,), torch.float32, dev) if self.any_grad else None
        self.garena = self.garena_full[: self.total] if self.any_grad else None
        self.gtail = self.garena_full[self.total:] if self.any_grad else None
        with torch.no_grad():
            for p in ordered:
                o, n = self.offsets[id(p)], p.numel()
                view = self.arena[o:o + n].view(p.shape)
                view.copy_(p.data)  # one-time setup copy
                p.data = view
        self.params = ordered
        self.bind_grads()
        # optional second CUDA stream for the weight-gradient GEMMs of the transformer: they are off the critical
        # path of the backward pass (nothing downstream reads them) and fill the tail of the dgrad kernels
        self.side_stream = None
        # called during the backward pass with a key of `grad_ranges` once every gradient of that slice has been issued
        self.grad_hook = None

        self.alias_params = []   # parameters of a real model that are views of this (padded shadow) model's
        self.dtype = None
        self.lin16 = None
        self.conv_w = {}     # layer index -> (wf, wd) in compute dtype (bf16 mode only)
        self.pos_w = None    # (norm, wp, wdp)
        self.dirty = True
        self.version_key = None
        self._hc_tables = None
        self._rel_tables = {}

    # ------------------------------------------------------------------ parameter / gradient views
    # views of the (never re-allocated) arenas are cached: an eager step asks for ~2500 of them
    def view(self, p):
        v = self._vcache.get(id(p))
        if v is None:
            o = self.offsets[id(p)]
            v = self._vcache[id(p)] = self.arena[o:o + p.numel()].view(p.shape)
        return v

    def gview(self, p):
        v = self._gcache.get(id(p))
        if v is None:
            o = self.offsets[id(p)]
            v = self._gcache[id(p)] = self.garena[o:o + p.numel()].view(p.shape)
        return v

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

    def rebind_grad_arena(self, storage):
        """Move the gradient arena into caller-provided storage of the same size (`total + GRAD_TAIL` floats) — the
        data-parallel step keeps it in symmetric memory so that ranks can reduce it over NVSwitch multicast."""
        assert self.garena_full is not None and storage.numel() == self.garena_full.numel() and storage.dtype == torch.float32
        with torch.no_grad():
            storage.copy_(self.garena_full)
        self.garena_full = storage
        self.garena = storage[: self.total]
        self.gtail = storage[self.total:]
        self._gcache = {}
        for p in self.params:
            if p.requires_grad:
                p.grad = self.gview(p)

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
        return sum(p._version for p in self.params) + sum(p._version for p in self.alias_params)

    def mark_dirty(self):
        self.dirty = True

    def refresh(self, dtype):
        """Rebuild the derived GEMM operands if any parameter changed (or the compute dtype did)."""
        key = self._vkey()
        if not self.dirty and self.dtype == dtype and key == self.version_key:
            return
        self.rebuild_derived(dtype, cast_lin=True)
        self.version_key = key

    def rebuild_derived(self, dtype, cast_lin=True):
        """bf16 copy of the Linear weights, tap-major conv weights, weight-normalised positional conv weights."""
        model = self.model
        dev = self.device
        bf = dtype == torch.bfloat16
        if bf:
            if self.lin16 is None:
                self.lin16 = torch.empty(self.lin_total, dtype=torch.bfloat16, device=dev)
                self._wcache = {}
                cast_lin = True
            if cast_lin:
                ops.cast(self.arena[: self.lin_total], self.lin16)
            for i, blk in enumerate(model.feature_extractor.conv_layers):
                if i == 0:
                    continue
                w = blk.conv.weight
                Cout, Cin, k = w.shape
                if Cin % 64 or Cout % 8:
                    raise DphShapeError(
                        "bf16 tensor-core mode needs conv input channels in multiples of 64 and output channels in "
                        "multiples of 8 (layer %d has %d -> %d); use compute dtype float32 for ragged pruned models"
                        % (i, Cin, Cout))
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

    def w16(self, p):
        """bf16 copy of a Linear weight (view into the bf16 arena)."""
        v = self._wcache.get(id(p))
        if v is None:
            o = self.offsets[id(p)]
            v = self._wcache[id(p)] = self.lin16[o:o + p.numel()].view(p.shape)
        return v

    def W(self, p, dtype):
        """GEMM operand copy of a Linear weight in the compute dtype."""
        return self.w16(p) if dtype == torch.bfloat16 else self.view(p)

    def wcat(self, ps, dtype):
        """Row-concatenation of consecutive Linear weights (same in_features) as one (sum out, in) matrix — valid
        because the arena lays them out back to back (used for the fused K/V/Q projection)."""
        o0 = self.offsets[id(ps[0])]
        o = o0
        for q in ps:
            if self.offsets[id(q)] != o or q.numel() % ALIGN:
                raise DphShapeError("parameters are not contiguous in the arena")
            o += q.numel()
        base = self.lin16 if dtype == torch.bfloat16 else self.arena
        return base[o0:o].view(-1, ps[0].shape[1])

    def vcat(self, ps):
        """Concatenation of consecutive fp32 vectors (biases) in the arena."""
        o0 = self.offsets[id(ps[0])]
        o = o0
        for q in ps:
            if self.offsets[id(q)] != o or q.numel() % ALIGN:
                raise DphShapeError("bias vectors are not contiguous in the arena")
            o += q.numel()
        return self.arena[o0:o]


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
_SHADOWS = {}   # id(model) -> padding.Shadow: ragged pruned models run on a zero-padded structural copy in bf16 mode


def _plan_of(model) -> Plan:
    plan = _PLANS.get(id(model))
    if plan is not None and plan.model is model and plan.still_bound() and len(plan.params) == sum(
            1 for _ in model.parameters()):
        return plan
    plan = Plan(model)
    _PLANS[id(model)] = plan
    return plan


def get_plan(model) -> Plan:
    from . import padding

    sh = _SHADOWS.get(id(model))
    if sh is not None and (sh.true is not model or not sh.aliased()):
        _SHADOWS.pop(id(model))
        _PLANS.pop(id(sh.model), None)
        sh = None
    if sh is None and model.compute_dtype == torch.bfloat16 and padding.is_ragged(model):
        _require_cuda(next(model.parameters()), "the model")
        sh = padding.Shadow(model)
        plan = _plan_of(sh.model)
        sh.alias()
        plan.alias_params = [p for p, _ in sh.pairs]
        _PLANS.pop(id(model), None)
        _SHADOWS[id(model)] = sh
    if sh is not None:
        sh.sync()
        return _plan_of(sh.model)
    return _plan_of(model)


def invalidate(model):
    _PLANS.pop(id(model), None)
    sh = _SHADOWS.pop(id(model), None)
    if sh is not None:
        _PLANS.pop(id(sh.model), None)


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
    if not _is_native(p0):
        return None
    plan = get_plan(model)
    tb = hc_tables(plan)
    if tb.nseg == 0:
        return None
    return _NumParams.apply(plan, *plan.la_params)


# ----------------------------------------------------------------------------------------------------------------
# Convolutional feature extractor (components.py:94-120, 158-185)
# ----------------------------------------------------------------------------------------------------------------
def _valid(ln):
    """Number of real columns of a LayerNorm whose width was zero-padded (see padding.py); 0 = all."""
    return int(getattr(ln, "valid_features", 0))


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


def conv_stack_fwd(plan: Plan, wave, masks, save: bool, A=None):
    """wave (B, T0) fp32 -> features (B, T, C) in the compute dtype. Returns (features, saved-state dict)."""
    model = plan.model
    fe = model.feature_extractor
    dt = model.compute_dtype
    dev = plan.device
    B, T0 = wave.shape
    wave = wave.contiguous()
    if A is None:
        A = FreshAlloc(dev)
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
                colmask = A("conv_last_colmask", (dummy.numel(),), torch.float32)
                ops.vec_mul(mask, dummy, colmask)
        ln = blk.layer_norm
        is_gn = isinstance(ln, torch.nn.GroupNorm)
        bias = plan.view(conv.bias) if conv.bias is not None else None
        rec = {"blk": blk, "i": i, "Tin": Tin, "Tout": Tout, "Cin": Cin, "Cout": Cout, "colmask": colmask,
               "has_mask": mask is not None, "x_in": x}
        y = A("conv%d_y" % i, (B, Tout, Cout), dt)
        if i == 0:
            if Cin != 1:
                raise DphShapeError("first conv layer must have one input channel")
            w = plan.view(conv.weight)
            if is_gn:
                if ln.num_groups != Cout:
                    raise DphShapeError("GroupNorm must have one group per channel")
                moments = A("conv0_moments", (B * 65,), torch.float64)
                mean = A("conv0_mean", (B * Cout,), torch.float32)
                rstd = A("conv0_rstd", (B * Cout,), torch.float32)
                ops.conv0_stats(wave, w, bias, moments, mean, rstd, B, Cout, k, s, Tout, ln.eps)
                ops.conv0_fwd(wave, w, bias, plan.view(ln.weight), plan.view(ln.bias), mean, rstd, colmask, y, B, Cout,
                              k, s, Tout, 1, 1)
                rec.update(mean=mean, rstd=rstd, moments=moments, mode="gn")
            elif ln is not None:
                z = A("conv%d_z" % i, (B, Tout, Cout), dt)
                ops.conv0_fwd(wave, w, bias, None, None, None, None, None, z, B, Cout, k, s, Tout, 0, 0)
                mean = A("conv%d_mean" % i, (B * Tout,), torch.float32)
                rstd = A("conv%d_rstd" % i, (B * Tout,), torch.float32)
                ops.layernorm_fwd(z, plan.view(ln.weight), plan.view(ln.bias), y, B * Tout, Cout, ln.eps, mean, rstd,
                                  colmask=colmask, act=1, d_valid=_valid(ln))
                rec.update(z=z, mean=mean, rstd=rstd, mode="ln")
            else:
                ops.conv0_fwd(wave, w, bias, None, None, None, None, colmask, y, B, Cout, k, s, Tout, 0, 1)
                rec.update(mode="none")
        else:
            fwd_w, _ = _conv_weight_views(plan, blk, i, dt)
            a = View(x, Tout, k * Cin, s * Cin, sb0=Tin * Cin)
            if ln is None:
                z = A("conv%d_z" % i, (B, Tout, Cout), dt) if save else None
                G.gemm(a, fwd_w, y, Tout, Cout, k * Cin, nb0=B, c_sb0=Tout * Cout, bias=bias, act=1, colmask=colmask,
                       c2=z)
                rec.update(z=z, mode="none")
            else:
                if is_gn:
                    raise DphShapeError("GroupNorm is only supported on the first conv layer (as in the reference)")
                z = A("conv%d_z" % i, (B, Tout, Cout), dt)
                G.gemm(a, fwd_w, z, Tout, Cout, k * Cin, nb0=B, c_sb0=Tout * Cout, bias=bias)
                mean = A("conv%d_mean" % i, (B * Tout,), torch.float32)
                rstd = A("conv%d_rstd" % i, (B * Tout,), torch.float32)
                ops.layernorm_fwd(z, plan.view(ln.weight), plan.view(ln.bias), y, B * Tout, Cout, ln.eps, mean, rstd,
                                  colmask=colmask, act=1, d_valid=_valid(ln))
                rec.update(z=z, mean=mean, rstd=rstd, mode="ln")
        rec["y"] = y
        st["layers"].append(rec)
        x = y
        Tin, Cin = Tout, Cout
    return x, st


# ----------------------------------------------------------------------------------------------------------------
# Buffer allocation: fresh tensors per call (autograd / module API) or a static, named workspace (training step:
# addresses never change, so the whole step can be captured in a CUDA graph)
# ----------------------------------------------------------------------------------------------------------------
class FreshAlloc:
    static = False

    def __init__(self, device):
        self.device = device

    def __call__(self, name, shape, dtype, zero=False):
        t = torch.empty(shape, dtype=dtype, device=self.device)
        if zero:
            ops.memset(t)
        return t


class StaticAlloc:
    """Named workspace with fixed addresses.  A buffer is keyed by its NAME only: asking for the same name with a
    different shape (the reference crops every batch to its shortest utterance, audio_dataset.py:285-363, so nearly
    every batch has its own (B, T0)) hands out a view of the same storage, which grows only when a larger size than
    ever before is requested — the workspace is bounded by the largest batch, not by the number of distinct shapes.
    `zero=True` buffers hold staging layouts whose padding is never written afterwards: they are cleared at creation
    and again whenever the requested shape (hence the position of the padding) changes."""
    static = True
    SLACK = 1.0   # growth factor applied when a buffer has to be re-allocated

    def __init__(self, device, prefix=""):
        self.device = device
        self.prefix = prefix
        self.bufs = {}    # name -> flat storage tensor
        self.shapes = {}  # name -> (shape, dtype) of the last view handed out
        self.views = {}   # name -> last view
        self.seen = set() # every (name, shape, dtype) ever handed out
        self.frozen = False  # set once a CUDA graph holds these addresses: shapes may no longer change
        self.generation = 0  # bumped whenever a storage is re-allocated: graphs captured before hold stale addresses
        # shape-keyed graph cache (trainer): several captured graphs of different shapes share this storage, so the
        # clearing of a `zero=True` staging buffer has to be part of every graph instead of happening once
        self.always_zero = False

    def __call__(self, name, shape, dtype, zero=False):
        name = self.prefix + name
        shape = tuple(int(s) for s in shape)
        key = (shape, dtype)
        if self.shapes.get(name) == key:
            if zero and self.always_zero:
                ops.memset(self.views[name])
            return self.views[name]
        if self.frozen and (name, key) not in self.seen:
            raise _lib.DphError(
                "workspace buffer %r was captured into a CUDA graph as %s and is now requested as %s: a captured step "
                "runs fixed shapes only (use run_step for variable-length batches)" % (name, self.shapes.get(name), key))
        n = 1
        for s in shape:
            n *= s
        store = self.bufs.get(name)
        if store is None or store.dtype != dtype or store.numel() < n:
            if store is not None:
                if store.is_cuda and torch.cuda.is_current_stream_capturing():
                    raise _lib.DphError("workspace buffer %r would have to grow during CUDA graph capture" % name)
                if store.is_cuda:
                    # kernels on side streams may still read the old storage; growth is rare (a new maximum size)
                    torch.cuda.synchronize(self.device)
            store = torch.empty(max(int(n * self.SLACK), n, 1), dtype=dtype, device=self.device)
            self.bufs[name] = store
            self.generation += 1
        t = store[:n].view(shape)
        if zero:
            ops.memset(t)
        self.shapes[name] = key
        self.views[name] = t
        self.seen.add((name, key))
        return t

    def nbytes(self):
        return sum(t.numel() * t.element_size() for t in self.bufs.values())


def _p(module, train):
    """dropout probability of an nn.Dropout in the current mode"""
    return float(module.p) if train else 0.0


def _seed(p):
    return ops.next_seed() if p > 0.0 else 0


# ----------------------------------------------------------------------------------------------------------------
# Convolutional feature extractor — backward
# ----------------------------------------------------------------------------------------------------------------
def conv_stack_bwd(plan: Plan, st, dfeat, dmasks, A):
    """dfeat: gradient wrt the (B, T, C) extractor output. Accumulates parameter gradients into the gradient arena
    and mask gradients into `dmasks` (flat, log_alpha layout)."""
    model = plan.model
    fe = model.feature_extractor
    dt = model.compute_dtype
    B = st["B"]
    wave = st["wave"]
    dy = dfeat
    pre_dzp = None   # set by layer i + 1 when its dgrad GEMMs already produced this layer's dz (fused epilogue backward)
    for rec in reversed(st["layers"]):
        blk, i = rec["blk"], rec["i"]
        conv = blk.conv
        Cout, Cin, k, s = rec["Cout"], rec["Cin"], blk.kernel_size, blk.stride
        Tin, Tout = rec["Tin"], rec["Tout"]
        ln = blk.layer_norm
        colmask = rec["colmask"]
        dmask = None
        if rec["has_mask"]:
            o = plan.offsets[id(blk.hard_concrete.log_alpha)] - plan.la_off
            dmask = dmasks[o:o + Cout]
        # the last layer's colmask is mask * dummy_weight: d mask = d colmask * dummy
        dm_tmp = None
        if dmask is not None and i == len(fe.conv_layers) - 1:
            dm_tmp = A("conv_dm_tmp", (Cout,), torch.float32)
            ops.memset(dm_tmp)
        dm_out = dm_tmp if dm_tmp is not None else dmask
        gbias = plan.gview(conv.bias) if (conv.bias is not None and conv.bias.requires_grad) else None
        bias = plan.view(conv.bias) if conv.bias is not None else None
        wreq = conv.weight.requires_grad
        if i == 0:
            w = plan.view(conv.weight)
            scratch = A("conv0_scratch", (16 * B * Cout,), torch.float32)
            if rec["mode"] == "gn":
                ops.conv0_bwd(wave, w, bias, plan.view(ln.weight), plan.view(ln.bias), rec["mean"], rec["rstd"],
                              rec["moments"], colmask, dy, scratch, plan.gview(conv.weight), gbias, plan.gview(ln.weight), plan.gview(ln.bias),
                              dm_out, B, Cout, k, s, Tout, 1, 1)
            elif rec["mode"] == "ln":
                dz = A("conv0_dz", (B, Tout, Cout), dt)
                ops.layernorm_bwd(dy, rec["z"], plan.view(ln.weight), plan.view(ln.bias), rec["mean"], rec["rstd"], dz,
                                  plan.gview(ln.weight), plan.gview(ln.bias), B * Tout, Cout, colmask=colmask,
                                  dmask=dm_out, act=1, d_valid=_valid(ln))
                ops.conv0_bwd(wave, w, bias, None, None, None, None, None, None, dz, scratch, plan.gview(conv.weight), gbias,
                              None, None, None, B, Cout, k, s, Tout, 0, 0)
            else:
                ops.conv0_bwd(wave, w, bias, None, None, None, None, None, colmask, dy, scratch, plan.gview(conv.weight),
                              gbias, None, None, dm_out, B, Cout, k, s, Tout, 0, 1)
        else:
            # dz in a per-utterance zero-padded row layout so the dgrad GEMM can read rows t'-j out of range
            Jmax = (k + s - 1) // s
            padr = Jmax + 1
            rows_pb = Tout + 2 * padr
            dzp = A("conv%d_dzp" % i, (B, rows_pb, Cout), dt, zero=True) if pre_dzp is None else pre_dzp
            if pre_dzp is not None:
                pre_dzp = None   # dz, the mask gradient and the bias gradient of this layer are already in place
            elif rec["mode"] == "ln":
                ops.layernorm_bwd(dy, rec["z"], plan.view(ln.weight), plan.view(ln.bias), rec["mean"], rec["rstd"], dzp,
                                  plan.gview(ln.weight), plan.gview(ln.bias), B * Tout, Cout, colmask=colmask,
                                  dmask=dm_out, act=1, rows_per_batch=Tout, out_bs=rows_pb * Cout, out_row_off=padr,
                                  d_valid=_valid(ln))
                if gbias is not None:
                    # bias gradient = column sum of dz
                    for b in range(B):
                        ops.epilogue_bwd(dzp[b, padr:padr + Tout], Tout, Cout, dbias=gbias)
            else:
                ops.epilogue_bwd(dy, B * Tout, Cout, z=rec["z"], colmask=colmask, dout=dzp, dmask=dm_out, dbias=gbias,
                                 act=1, rows_per_batch=Tout, out_bs=rows_pb * Cout, out_row_off=padr)
            x_in = rec["x_in"]
            # ---- weight gradient: dW[co][tap][ci] = sum_{b,t} dz[b,t,co] * x[b, s t + tap, ci]
            if wreq:
                wg = A("conv_wg_scratch", (Cout, k * Cin), torch.float32)
                ops.memset(wg)
                a = View(dzp, Cout, Tout, 1, cs=Cout, sb0=rows_pb * Cout, offset=padr * Cout)
                bv = View(x_in, k * Cin, Tout, 1, cs=s * Cin, sb0=Tin * Cin)
                G.gemm(a, bv, wg, Cout, k * Cin, Tout, nb0=B, c_sb0=0, atomic=True, ksplit=-1)
                ops.conv_wgrad_unpermute(wg, plan.gview(conv.weight), Cout, Cin, k)
            # ---- input gradient, one GEMM per output phase (u = s t' + phase)
            _, phases = _conv_weight_views(plan, blk, i, dt)
            # the layer below, when it is a plain conv -> GELU -> mask block: its epilogue backward rides in this
            # layer's dgrad GEMMs, which then write d z of the layer below straight into ITS padded row layout
            below = st["layers"][i - 1] if i >= 2 else None
            fuse = None
            if below is not None and FUSED_ACT_BWD and below["mode"] == "none" and below["z"] is not None and k >= s:
                kb_, sb_ = below["blk"].kernel_size, below["blk"].stride
                padr_b = (kb_ + sb_ - 1) // sb_ + 1
                rows_pb_b = Tin + 2 * padr_b
                dzp_b = A("conv%d_dzp" % (i - 1), (B, rows_pb_b, Cin), dt, zero=True)
                if G.act_bwd_ok(dzp_b, below["z"], Cin, s * Cin, s * Cin, (rows_pb_b * Cin, Tin * Cin),
                                [(padr_b + ph) * Cin for ph in range(s)] + [ph * Cin for ph in range(s)]):
                    bconv = below["blk"].conv
                    fuse = dict(dzp=dzp_b, padr=padr_b, rows_pb=rows_pb_b, z=below["z"], colmask=below["colmask"],
                                dmask=(dmasks[plan.offsets[id(below["blk"].hard_concrete.log_alpha)] - plan.la_off:][:Cin]
                                       if below["has_mask"] else None),
                                dbias=(plan.gview(bconv.bias) if (bconv.bias is not None and bconv.bias.requires_grad)
                                       else None))
            dx = A("conv%d_dx" % i, (B, Tin, Cin), dt) if fuse is None else None
            covered = set()
            for ph, J, wv in phases:
                rows = (Tin - ph + s - 1) // s
                if rows <= 0:
                    continue
                a = View(dzp, rows, J * Cout, Cout, kin=Cout, kxs=Cout, sb0=rows_pb * Cout,
                         offset=(padr - (J - 1)) * Cout)
                if fuse is None:
                    G.gemm(a, wv, dx, rows, Cin, J * Cout, nb0=B, ldc=s * Cin, c_sb0=Tin * Cin, c_offset=ph * Cin)
                else:
                    G.gemm(a, wv, fuse["dzp"], rows, Cin, J * Cout, nb0=B, ldc=s * Cin, c_sb0=fuse["rows_pb"] * Cin,
                           c_offset=(fuse["padr"] + ph) * Cin, act=1, colmask=fuse["colmask"], act_bwd=True, z=fuse["z"],
                           ldz=s * Cin, z_sb0=Tin * Cin, z_offset=ph * Cin, dbias=fuse["dbias"], dmask=fuse["dmask"])
                covered.add(ph)
            if fuse is not None:
                pre_dzp = fuse["dzp"]
            for ph in range(s):
                if ph not in covered and dx is not None:  # stride > kernel: some input rows get no gradient
                    rows = (Tin - ph + s - 1) // s
                    for b in range(B):
                        ops.memset2d(dx, (b * Tin + ph) * Cin, s * Cin, Cin, rows)
            dy = dx
        if dm_tmp is not None:
            ops.vec_mul(dm_tmp, plan.view(fe.dummy_weight), dmask, 1.0, 1.0)
        if i == 1 and plan.grad_hook is not None:
            plan.grad_hook("conv_hi")   # conv layers n-1 .. 1 are done; only layer 0 is left


# ----------------------------------------------------------------------------------------------------------------
# Encoder forward (components.py:263-274, 319-333, 379-436, 604-659, 726-748, 814-857, 885-927, 968-1003)
# ----------------------------------------------------------------------------------------------------------------
def _a8(n):
    return (n + 7) // 8 * 8


# bf16 / head_dim 64: attention runs as the fused tcgen05 kernel (scores never reach HBM); set False to force the
# unfused GEMM + softmax path (used by tests to cross-check the two)
FUSED_ATTENTION = True
# bf16: the backward of GELU / column-mask / dropout epilogues runs inside the dgrad GEMM that produces its input
# (csrc/gemm_kernel.cuh: epilogue_chunk_bwd); False restores the separate elementwise pass (tests cross-check the two)
FUSED_ACT_BWD = True


def _mask_of(masks, hc):
    return masks.get(id(hc)) if hc is not None else None


def _rel_table(plan, att0, T):
    """Bucket index of every relative offset j - i in [-(T-1), T-1] — the integer part of
    WavLMSelfAttention.compute_bias / _relative_positions_bucket (components.py:546-600), built on the host."""
    key = (T, att0.num_buckets, att0.max_distance)
    tb = plan._rel_tables.get(key)
    if tb is None:
        rel = torch.arange(-(T - 1), T, dtype=torch.long)
        nb = att0.num_buckets // 2
        buckets = (rel > 0).to(torch.long) * nb
        rp = rel.abs()
        max_exact = nb // 2
        is_small = rp < max_exact
        large = max_exact + (torch.log(rp.float() / max_exact) / math.log(att0.max_distance / max_exact)
                             * (nb - max_exact)).to(torch.long)
        large = torch.min(large, torch.full_like(large, nb - 1))
        buckets = buckets + torch.where(is_small, rp, large)
        tb = buckets.to(torch.int32).to(plan.device)
        plan._rel_tables[key] = tb
    return tb


def _head_index(plan, att):
    key = ("hidx", id(att))
    t = plan._rel_tables.get(key)
    if t is None:
        t = torch.tensor(list(att.remaining_heads), dtype=torch.int32).to(plan.device)
        plan._rel_tables[key] = t
    return t


def _attention_fwd(plan, att, xin, resid, B, T, lengths, masks, A, tag, train, p_drop, wl, save=True):
    """xin: attention input (M, D); resid: residual added after dropout. Returns (resid + dropout(attn)), record."""
    dt = xin.dtype
    D, H, Dh = att.embed_dim, att.num_heads, att.head_dim
    inner = H * Dh
    M = B * T
    Tp = _a8(T)
    rec = {"att": att, "xin": xin, "H": H}
    Wkvq = plan.wcat([att.k_proj.weight, att.v_proj.weight, att.q_proj.weight], dt)
    bkvq = plan.vcat([att.k_proj.bias, att.v_proj.bias, att.q_proj.bias])
    qkv = A(tag + "qkv", (M, 3 * inner), dt)
    G.gemm(G.mat(xin), G.mat(Wkvq), qkv, M, 3 * inner, D, bias=bkvq)
    s_t, s_b = 3 * inner, T * 3 * inner
    wavlm = None
    if wl is not None:
        # gated relative position bias; the gate is a function of the attention input (components.py:636-644)
        Htot = att.total_num_heads
        gate = ga = gb = None
        if att.gru_rel_pos:
            gate = A(tag + "gate", (B, Htot, T), torch.float32)
            ga = A(tag + "ga", (B, Htot, T), torch.float32)
            gb = A(tag + "gb", (B, Htot, T), torch.float32)
            ops.wavlm_gate_fwd(xin, plan.view(att.gru_rel_pos_linear.weight), plan.view(att.gru_rel_pos_linear.bias),
                               plan.view(att.gru_rel_pos_const), gate, ga, gb, B, T, Htot, D // Htot)
        wavlm = (wl["rel_table"], wl["emb"], gate, _head_index(plan, att), Htot)
        rec.update(ga=ga, gb=gb, wavlm=wavlm)
    p_att = _p(att.dropout, train)
    seed_att = _seed(p_att)
    hmask = _mask_of(masks, att.hard_concrete_for_heads)
    lmask = _mask_of(masks, att.hard_concrete_for_layer)
    ctx = A(tag + "ctx", (M, inner), dt)
    ctx_un = A(tag + "ctx_un", (M, inner), dt) if (hmask is not None and save) else None
    fused = FUSED_ATTENTION and dt == torch.bfloat16 and Dh == 64 and T <= 1024
    P = Pd = lse = None
    if fused:
        # S and P never leave the SM: tcgen05 QK^T -> online softmax -> PV  (csrc/attention.cu)
        lse = A(tag + "lse", (B, H, T), torch.float32)
        ops.attn_fwd(qkv, ctx, ctx_un, lse, B, H, T, att.scaling, lengths, p_att, seed_att, wavlm, hmask)
    else:
        S = A("att_S", (B, H, T, Tp), torch.float32)
        q = View(qkv, T, Dh, s_t, sb0=Dh, sb1=s_b, offset=2 * inner)
        k = View(qkv, T, Dh, s_t, sb0=Dh, sb1=s_b, offset=0)
        G.gemm(q, k, S, T, T, Dh, nb0=H, nb1=B, ldc=Tp, c_sb0=T * Tp, c_sb1=H * T * Tp, alpha=att.scaling)
        P = A(tag + "P", (B, H, T, Tp), dt)
        Pd = A(tag + "Pd", (B, H, T, Tp), dt) if p_att > 0 else None
        ops.softmax_fwd(S, Tp, P, Pd, Tp, lengths, B, H, T, p_att, seed_att, wavlm)
        pv = View(Pd if Pd is not None else P, T, T, Tp, sb0=T * Tp, sb1=H * T * Tp)
        v = View(qkv, Dh, T, 1, cs=s_t, sb0=Dh, sb1=s_b, offset=inner)
        G.gemm(pv, v, ctx, T, Dh, T, nb0=H, nb1=B, ldc=inner, c_sb0=Dh, c_sb1=T * inner, scalar_mask=hmask,
               scalar_sb0=1 if hmask is not None else 0, c2=ctx_un)
    out = A(tag + "attn_out", (M, D), dt)
    pre = A(tag + "attn_pre", (M, D), dt) if (lmask is not None and save) else None
    seed_o = _seed(p_drop)
    G.gemm(G.mat(ctx), G.mat(plan.W(att.out_proj.weight, dt)), out, M, D, inner, bias=plan.view(att.out_proj.bias),
           scalar_mask=lmask, drop_p=p_drop, seed=seed_o, residual=resid, c2=pre)
    rec.update(qkv=qkv, P=P, Pd=Pd, ctx=ctx, ctx_un=ctx_un, pre=pre, hmask=hmask, lmask=lmask, p_att=p_att,
               seed_att=seed_att, p_drop=p_drop, seed_o=seed_o, lse=lse, lengths=lengths)
    return out, rec


def _ffn_fwd(plan, ff, xin, resid, M, masks, A, tag, train, save=True):
    dt = xin.dtype
    D = ff.intermediate_dense.in_features
    F = ff.intermediate_dense.out_features
    imask = _mask_of(masks, ff.hard_concrete_for_intermediate)
    lmask = _mask_of(masks, ff.hard_concrete_for_layer)
    p_i, p_o = _p(ff.intermediate_dropout, train), _p(ff.output_dropout, train)
    seed_i, seed_o = _seed(p_i), _seed(p_o)
    h = A(tag + "ff_h", (M, F), dt)
    zpre = A(tag + "ff_z", (M, F), dt) if save else None   # pre-activation copy: the backward's input only
    G.gemm(G.mat(xin), G.mat(plan.W(ff.intermediate_dense.weight, dt)), h, M, F, D,
           bias=plan.view(ff.intermediate_dense.bias), act=1, colmask=imask, drop_p=p_i, seed=seed_i, c2=zpre)
    out = A(tag + "ff_out", (M, D), dt)
    pre = A(tag + "ff_pre", (M, D), dt) if (lmask is not None and save) else None
    G.gemm(G.mat(h), G.mat(plan.W(ff.output_dense.weight, dt)), out, M, D, F, bias=plan.view(ff.output_dense.bias),
           scalar_mask=lmask, drop_p=p_o, seed=seed_o, residual=resid, c2=pre)
    rec = {"ff": ff, "xin": xin, "h": h, "zpre": zpre, "pre": pre, "imask": imask, "lmask": lmask, "p_i": p_i,
           "seed_i": seed_i, "p_o": p_o, "seed_o": seed_o, "F": F}
    return out, rec


def _ln_fwd(plan, ln, x, rows, D, A, tag, drop_p=0.0):
    y = A(tag + "y", (rows, D), x.dtype)
    mean = A(tag + "mean", (rows,), torch.float32)
    rstd = A(tag + "rstd", (rows,), torch.float32)
    seed = _seed(drop_p)
    ops.layernorm_fwd(x, plan.view(ln.weight), plan.view(ln.bias), y, rows, D, ln.eps, mean, rstd, drop_p=drop_p,
                      seed=seed, d_valid=_valid(ln))
    return y, {"ln": ln, "x": x, "mean": mean, "rstd": rstd, "drop_p": drop_p, "seed": seed}


def _posconv_geometry(pc, dt):
    conv = pc.conv
    D = pc.embed_dim
    K = pc.kernel_size
    Gn = conv.groups
    Cg = D // Gn
    Cgp = _align_to(Cg, 64) if dt == torch.bfloat16 else Cg
    return D, K, Gn, Cg, Cgp


def _posconv_fwd(plan, pc, h0, B, T, A, tag, save=True):
    """x + GELU(conv(x)) with the weight-normalised grouped conv as one batched implicit GEMM over a zero-padded,
    group-padded staging copy of x (rows: front + b * (T + pad) + t)."""
    dt = h0.dtype
    D, K, Gn, Cg, Cgp = _posconv_geometry(pc, dt)
    pad = K  # zero rows between utterances (>= K - 1 so windows never see a neighbour)
    front = K // 2
    rows = front + B * (T + pad) + K
    Wd = Gn * Cgp
    xp = A(tag + "pos_xp", (rows, Wd), dt, zero=True)
    ops.posconv_pad(h0, xp, B, T, Gn, Cg, Cgp, front, pad)
    _, wp, _, _ = plan.pos_w
    zpc = A(tag + "pos_z", (B * T, D), dt) if save else None
    xpre = A(tag + "pos_out", (B * T, D), dt)
    # output frame t reads staging rows (front + b (T+pad) + t) - K/2 + tap
    if ops.posconv_tc_ok(dt, Cg, Cgp, K):
        ops.posconv_tc(xp, wp, plan.view(pc.conv.bias), h0, xpre, zpc, B, T, D, Gn, Cg, Cgp, K, T + pad, front - K // 2, 1)
    else:
        a = View(xp, T, K * Cgp, Wd, kin=Cgp, kxs=Wd, sb0=Cgp, sb1=(T + pad) * Wd, offset=(front - K // 2) * Wd)
        bv = View(wp, Cg, K * Cgp, K * Cgp, kin=Cgp, kxs=Cgp, sb0=Cg * K * Cgp)
        G.gemm(a, bv, xpre, T, Cg, K * Cgp, nb0=Gn, nb1=B, ldc=D, c_sb0=Cg, c_sb1=T * D, bias=plan.view(pc.conv.bias),
               vec_sb0=Cg, act=1, residual=h0, ldr=D, r_sb0=Cg, r_sb1=T * D, c2=zpc)
    return xpre, {"pc": pc, "xp": xp, "zpc": zpc, "front": front, "pad": pad, "rows": rows}


def encoder_fwd(plan: Plan, feat, lengths, masks, A, train, num_layers=None, final_norm=False, layer_keep=None,
                save=True):
    """feat (B, T, C) compute dtype -> list of hidden states [(B, T, D)] (index 0 = encoder input after padding
    zeroing), plus the record needed by `encoder_bwd`."""
    model = plan.model
    enc = model.encoder
    tr = enc.transformer
    fp = enc.feature_projection
    dt = feat.dtype
    B, T, C = feat.shape
    M = B * T
    D = fp.projection.out_features
    st = {"B": B, "T": T, "feat": feat, "lengths": lengths}
    # ---- feature projection: LN -> Linear -> dropout, padded frames zeroed
    xn, rec_ln = _ln_fwd(plan, fp.layer_norm, feat.view(M, C), M, C, A, "fp_ln_")
    p_in = _p(fp.dropout, train)
    seed_in = _seed(p_in)
    h0 = A("h0", (M, D), dt)
    G.gemm(G.mat(xn), G.mat(plan.W(fp.projection.weight, dt)), h0, M, D, C, bias=plan.view(fp.projection.bias),
           drop_p=p_in, seed=seed_in)
    if lengths is not None:
        ops.pad_rows_zero(h0, lengths, B, T, D)
    st["fp"] = {"ln": rec_ln, "xn": xn, "p_in": p_in, "seed_in": seed_in}
    hiddens = [h0]
    # ---- positional conv + (Base) LayerNorm + dropout
    xpre, rec_pc = _posconv_fwd(plan, tr.pos_conv_embed, h0, B, T, A, "", save)
    st["pc"] = rec_pc
    p_tr = _p(tr.dropout, train)
    if tr.layer_norm_first:
        x, rec = _ln_fwd(plan, tr.layer_norm, xpre, M, D, A, "tr_ln_", drop_p=p_tr)
        st["tr_ln"] = rec
    elif p_tr > 0:
        seed = _seed(p_tr)
        x = A("tr_drop", (M, D), dt)
        ops.epilogue_bwd(xpre, M, D, dout=x, drop_p=p_tr, seed=seed)  # plain dropout (no act / masks)
        st["tr_drop"] = {"p": p_tr, "seed": seed}
    else:
        x = xpre
    # ---- layers
    wl = None
    st["layers"] = []
    for li, layer in enumerate(tr.layers):
        if layer_keep is not None and not layer_keep[li]:
            st["layers"].append(None)
            continue
        tag = "L%d_" % li
        att, ff = layer.attention, layer.feed_forward
        rec = {"layer": layer, "x_in": x}
        p_drop = _p(layer.dropout, train)
        if att is not None:
            if getattr(att, "has_position_bias", False) and wl is None:
                att0 = att if att.rel_attn_embed is not None else None
                if att0 is None:
                    raise DphShapeError("WavLM attention without a relative position embedding in the first layer")
                wl = {"rel_table": _rel_table(plan, att0, T), "emb": plan.view(att0.rel_attn_embed.weight),
                      "att0": att0}
            wl_here = wl if getattr(att, "has_position_bias", False) else None
            if layer.layer_norm_first:
                xa, rec["ln1"] = _ln_fwd(plan, layer.layer_norm, x, M, D, A, tag + "ln1_")
            else:
                xa = x
            x1, rec["att"] = _attention_fwd(plan, att, xa, x, B, T, lengths, masks, A, tag, train, p_drop, wl_here, save)
        else:
            x1 = x
        if layer.layer_norm_first:
            if ff is not None:
                xf, rec["ln2"] = _ln_fwd(plan, layer.final_layer_norm, x1, M, D, A, tag + "ln2_")
                x2, rec["ff"] = _ffn_fwd(plan, ff, xf, x1, M, masks, A, tag, train, save)
            else:
                x2 = x1
        else:
            xl, rec["ln1"] = _ln_fwd(plan, layer.layer_norm, x1, M, D, A, tag + "ln1_")
            if ff is not None:
                xs, rec["ff"] = _ffn_fwd(plan, ff, xl, xl, M, masks, A, tag, train, save)
            else:
                xs = xl
            x2, rec["ln2"] = _ln_fwd(plan, layer.final_layer_norm, xs, M, D, A, tag + "ln2_")
        st["layers"].append(rec)
        x = x2
        hiddens.append(x)
        if num_layers is not None and len(hiddens) - 1 >= num_layers:
            break
    if final_norm and not tr.layer_norm_first:
        x, st["final_ln"] = _ln_fwd(plan, tr.layer_norm, x, M, D, A, "final_ln_")
        hiddens.append(x)
    st["wl"] = wl
    st["n_hidden"] = len(hiddens)
    return [h.view(B, T, D) for h in hiddens], st


# ----------------------------------------------------------------------------------------------------------------
# Encoder backward
# ----------------------------------------------------------------------------------------------------------------
def _ksplit(Mo, No, K):
    """Split-K factor for weight-gradient GEMMs (small outputs, long reductions): -1 lets the library pick it from
    the tile shape it chooses (bf16 tensor-core path); the fp32 path ignores it."""
    return -1


class _on_side:
    """Run the enclosed launches on the plan's side stream (after everything issued so far on the current stream).
    Inputs must stay untouched until the streams are joined again: every buffer a weight-gradient GEMM reads is a
    per-layer static buffer that is written once per step."""

    def __init__(self, plan):
        self.side = plan.side_stream
        self.ctx = None

    def __enter__(self):
        if self.side is not None:
            self.side.wait_stream(torch.cuda.current_stream())
            self.ctx = torch.cuda.stream(self.side)
            self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
        return False


def join_side(plan):
    """current stream waits for the side stream (end of the backward pass)"""
    if plan.side_stream is not None:
        torch.cuda.current_stream().wait_stream(plan.side_stream)


def _wgrad(plan, dy2d, x2d, param, Mo, No, Krows):
    """grad(param)[o, i] += sum_m dy[m, o] * x[m, i]   (fp32, atomically accumulated into the gradient arena)"""
    if not param.requires_grad:
        return
    with _on_side(plan):
        G.gemm(G.mat_t(dy2d), G.mat_t(x2d), plan.gview(param), Mo, No, Krows, atomic=True,
               ksplit=_ksplit(Mo, No, Krows))


def _gslot(plan, dmasks, hc):
    """slice of the flat mask-gradient buffer belonging to gate module `hc`"""
    if hc is None:
        return None
    o = plan.offsets[id(hc.log_alpha)] - plan.la_off
    return dmasks[o:o + hc.n_in]


def _gb(plan, p):
    return plan.gview(p) if (p is not None and p.requires_grad) else None


def _ln_bwd(plan, rec, dy, dy2, A, tag):
    ln = rec["ln"]
    x = rec["x"]
    rows, D = x.shape[0], x.shape[1]
    dx = A(tag + "dx", (rows, D), x.dtype)
    ops.layernorm_bwd(dy, x, plan.view(ln.weight), plan.view(ln.bias), rec["mean"], rec["rstd"], dx, _gb(plan, ln.weight),
                      _gb(plan, ln.bias), rows, D, dy2=dy2, drop_p=rec["drop_p"], seed=rec["seed"], d_valid=_valid(ln))
    return dx


def _out_proj_bwd(plan, dout, pre, lmask, hc_layer, bias_p, p_drop, seed, dmasks, A, tag, M, D):
    """Backward of  o = dropout(lmask * (lin + bias)) : returns d lin, accumulates d bias / d layer gate."""
    if p_drop > 0 or lmask is not None:
        g = A(tag + "g", (M, D), dout.dtype)
        ops.epilogue_bwd(dout, M, D, z=pre, scalar=lmask, dout=g, dbias=_gb(plan, bias_p),
                         dscalar=_gslot(plan, dmasks, hc_layer) if lmask is not None else None, drop_p=p_drop, seed=seed)
        return g
    if _gb(plan, bias_p) is not None:
        ops.epilogue_bwd(dout, M, D, dbias=_gb(plan, bias_p))
    return dout


def _ffn_bwd(plan, rec, dout, dresid, dmasks, A, tag):
    ff = rec["ff"]
    dt = dout.dtype
    M, D = dout.shape
    F = rec["F"]
    g = _out_proj_bwd(plan, dout, rec["pre"], rec["lmask"], ff.hard_concrete_for_layer, ff.output_dense.bias,
                      rec["p_o"], rec["seed_o"], dmasks, A, tag + "ffo_", M, D)
    dz = A(tag + "ff_dz", (M, F), dt)
    dm_i = _gslot(plan, dmasks, ff.hard_concrete_for_intermediate) if rec["imask"] is not None else None
    if FUSED_ACT_BWD and G.act_bwd_ok(dz, rec["zpre"], F, F, F):
        # d hidden = g W2 and the backward of  h = dropout(mask * GELU(z))  in ONE launch: the dgrad GEMM's epilogue
        # applies GELU'(z) * mask * dropout and reduces the bias / mask gradients (no (M, F) round trip through HBM)
        G.gemm(G.mat(g), G.mat_t(plan.W(ff.output_dense.weight, dt)), dz, M, F, D, act=1, colmask=rec["imask"],
               drop_p=rec["p_i"], seed=rec["seed_i"], act_bwd=True, z=rec["zpre"], dmask=dm_i,
               dbias=_gb(plan, ff.intermediate_dense.bias))
        _wgrad(plan, g, rec["h"], ff.output_dense.weight, D, F, M)
    else:
        dh = A(tag + "ff_dh", (M, F), dt)
        G.gemm(G.mat(g), G.mat_t(plan.W(ff.output_dense.weight, dt)), dh, M, F, D)
        _wgrad(plan, g, rec["h"], ff.output_dense.weight, D, F, M)
        ops.epilogue_bwd(dh, M, F, z=rec["zpre"], colmask=rec["imask"], dout=dz, dmask=dm_i,
                         dbias=_gb(plan, ff.intermediate_dense.bias), act=1, drop_p=rec["p_i"], seed=rec["seed_i"])
    dx = A(tag + "ff_dx", (M, D), dt)
    G.gemm(G.mat(dz), G.mat_t(plan.W(ff.intermediate_dense.weight, dt)), dx, M, D, F, residual=dresid)
    _wgrad(plan, dz, rec["xin"], ff.intermediate_dense.weight, F, D, M)
    return dx


def _attention_bwd(plan, rec, dout, dresid, B, T, dmasks, A, tag, wl):
    att = rec["att"]
    dt = dout.dtype
    M, D = dout.shape
    H, Dh = rec["H"], att.head_dim
    inner = H * Dh
    Tp = _a8(T)
    s_t, s_b = 3 * inner, T * 3 * inner
    g = _out_proj_bwd(plan, dout, rec["pre"], rec["lmask"], att.hard_concrete_for_layer, att.out_proj.bias,
                      rec["p_drop"], rec["seed_o"], dmasks, A, tag + "ao_", M, D)
    dctx = A(tag + "dctx", (M, inner), dt)
    G.gemm(G.mat(g), G.mat_t(plan.W(att.out_proj.weight, dt)), dctx, M, inner, D)
    _wgrad(plan, g, rec["ctx"], att.out_proj.weight, D, inner, M)
    qkv = rec["qkv"]
    wavlm = rec.get("wavlm")
    dgate = None
    demb = None
    nbuckets = 0
    if wavlm is not None:
        att0 = wl["att0"]
        nbuckets = att0.num_buckets
        demb = _gb(plan, att0.rel_attn_embed.weight)
        if wavlm[2] is not None:
            dgate = A(tag + "dgate", (B, att.total_num_heads, T), torch.float32)
            ops.memset(dgate)
    dqkv = A(tag + "dqkv", (M, 3 * inner), dt)
    if rec["lse"] is not None:
        delta = A("att_delta", (B, H, T), torch.float32)
        dq32 = A("att_dq32_%d" % inner, (M, inner), torch.float32, zero=True)
        o_un = rec["ctx_un"] if rec["hmask"] is not None else rec["ctx"]
        ops.attn_bwd(qkv, dctx, o_un, rec["lse"], delta, dq32, dqkv, _gslot(plan, dmasks, att.hard_concrete_for_heads)
                     if rec["hmask"] is not None else None, B, H, T, att.scaling, rec["lengths"], rec["p_att"],
                     rec["seed_att"], wavlm, rec["hmask"], dgate, demb, nbuckets)
    else:
        if rec["hmask"] is not None:
            dctx2 = A(tag + "dctx2", (M, inner), dt)
            ops.group_mask_bwd(dctx, rec["ctx_un"], rec["hmask"], dctx2,
                               _gslot(plan, dmasks, att.hard_concrete_for_heads), M, inner, Dh)
            dctx = dctx2
        P, Pd = rec["P"], rec["Pd"] if rec["Pd"] is not None else rec["P"]
        # dPd = dctx V^T
        dP = A("att_dP", (B, H, T, Tp), torch.float32)
        a = View(dctx, T, Dh, inner, sb0=Dh, sb1=T * inner)
        vk = View(qkv, T, Dh, s_t, sb0=Dh, sb1=s_b, offset=inner)
        G.gemm(a, vk, dP, T, T, Dh, nb0=H, nb1=B, ldc=Tp, c_sb0=T * Tp, c_sb1=H * T * Tp)
        # dV = Pd^T dctx
        a = View(Pd, T, T, 1, cs=Tp, sb0=T * Tp, sb1=H * T * Tp)
        bv = View(dctx, Dh, T, 1, cs=inner, sb0=Dh, sb1=T * inner)
        G.gemm(a, bv, dqkv, T, Dh, T, nb0=H, nb1=B, ldc=s_t, c_sb0=Dh, c_sb1=s_b, c_offset=inner)
        # dS = softmax'(P) . dropout'(dPd)  (+ WavLM bias gradients)
        dS = A("att_dS", (B, H, T, Tp), dt)
        ops.softmax_bwd(P, dP, Tp, dS, Tp, B, H, T, rec["p_att"], rec["seed_att"], wavlm, dgate, demb, nbuckets)
        # dQ = scaling * dS K ; dK = scaling * dS^T Q
        a = View(dS, T, T, Tp, sb0=T * Tp, sb1=H * T * Tp)
        bv = View(qkv, Dh, T, 1, cs=s_t, sb0=Dh, sb1=s_b, offset=0)
        G.gemm(a, bv, dqkv, T, Dh, T, nb0=H, nb1=B, ldc=s_t, c_sb0=Dh, c_sb1=s_b, c_offset=2 * inner, alpha=att.scaling)
        a = View(dS, T, T, 1, cs=Tp, sb0=T * Tp, sb1=H * T * Tp)
        bv = View(qkv, Dh, T, 1, cs=s_t, sb0=Dh, sb1=s_b, offset=2 * inner)
        G.gemm(a, bv, dqkv, T, Dh, T, nb0=H, nb1=B, ldc=s_t, c_sb0=Dh, c_sb1=s_b, c_offset=0, alpha=att.scaling)
    # projections
    kvq_w = [att.k_proj.weight, att.v_proj.weight, att.q_proj.weight]
    kvq_b = [att.k_proj.bias, att.v_proj.bias, att.q_proj.bias]
    if kvq_b[0].requires_grad:
        o0 = plan.offsets[id(kvq_b[0])]
        ops.epilogue_bwd(dqkv, M, 3 * inner, dbias=plan.garena[o0:o0 + 3 * inner])
    xin = rec["xin"]
    if kvq_w[0].requires_grad:
        o0 = plan.offsets[id(kvq_w[0])]
        gw = plan.garena[o0:o0 + 3 * inner * D].view(3 * inner, D)
        with _on_side(plan):
            G.gemm(G.mat_t(dqkv), G.mat_t(xin), gw, 3 * inner, D, M, atomic=True, ksplit=_ksplit(3 * inner, D, M))
    dx = A(tag + "att_dx", (M, D), dt)
    G.gemm(G.mat(dqkv), G.mat_t(plan.wcat(kvq_w, dt)), dx, M, D, 3 * inner, residual=dresid)
    if dgate is not None:
        lin = att.gru_rel_pos_linear
        ops.wavlm_gate_bwd(xin, plan.view(lin.weight), plan.view(att.gru_rel_pos_const), rec["ga"], rec["gb"], dgate, dx,
                           plan.gview(lin.weight), plan.gview(lin.bias), plan.gview(att.gru_rel_pos_const), B, T,
                           att.total_num_heads, D // att.total_num_heads)
    return dx


def _acc(total, extra, n):
    """total += extra (same dtype, n elements); either may be None. Returns the combined buffer."""
    if extra is None:
        return total
    if total is None:
        return extra
    ops.axpy(total, extra, n)
    return total


def encoder_bwd(plan: Plan, st, dh, dmasks, A):
    """dh: list (len n_hidden) of gradients wrt the returned hidden states ((M, D) compute dtype, or None).
    Returns the gradient wrt the extractor features (B, T, C). NOTE: buffers in `dh` may be modified in place."""
    model = plan.model
    enc = model.encoder
    tr = enc.transformer
    fp = enc.feature_projection
    B, T = st["B"], st["T"]
    M = B * T
    feat = st["feat"]
    C = feat.shape[2]
    D = fp.projection.out_features
    dt = feat.dtype
    dh = list(dh)
    g = None  # gradient flowing back from later layers wrt the current layer output
    idx = st["n_hidden"] - 1
    if "final_ln" in st:
        g = _ln_bwd(plan, st["final_ln"], dh[idx], None, A, "final_ln_")
        idx -= 1
    n_run = sum(1 for r in st["layers"] if r is not None)
    assert idx == n_run
    wl = st.get("wl")
    for li in range(len(st["layers"]) - 1, -1, -1):
        rec = st["layers"][li]
        if rec is None:
            continue
        layer = rec["layer"]
        tag = "L%d_" % li
        d_out, d_head = g, dh[idx]
        idx -= 1
        if d_out is None and d_head is None:
            g = None
            # nothing downstream depends on this layer's output (e.g. num_layers cut) -> no gradient
            continue
        if layer.layer_norm_first:
            d2 = _acc(d_out, d_head, M * D)
            if "ff" in rec:
                dxf = _ffn_bwd(plan, rec["ff"], d2, None, dmasks, A, tag)
                dx1 = _ln_bwd(plan, rec["ln2"], dxf, None, A, tag + "ln2_")
                ops.axpy(dx1, d2, M * D)
            else:
                dx1 = d2
            if "att" in rec:
                dxa = _attention_bwd(plan, rec["att"], dx1, None, B, T, dmasks, A, tag, wl)
                g = _ln_bwd(plan, rec["ln1"], dxa, None, A, tag + "ln1_")
                ops.axpy(g, dx1, M * D)
            else:
                g = dx1
        else:
            if d_out is None:
                d_out, d_head = d_head, None
            ds = _ln_bwd(plan, rec["ln2"], d_out, d_head, A, tag + "ln2_")
            dxl = _ffn_bwd(plan, rec["ff"], ds, ds, dmasks, A, tag) if "ff" in rec else ds
            dx1 = _ln_bwd(plan, rec["ln1"], dxl, None, A, tag + "ln1_")
            g = _attention_bwd(plan, rec["att"], dx1, dx1, B, T, dmasks, A, tag, wl) if "att" in rec else dx1
        if plan.grad_hook is not None:
            plan.grad_hook(li)  # every weight gradient of transformer layer li has been issued (data-parallel overlap)
    assert idx == 0
    d_h0_head = dh[0]
    if g is not None:
        # ---- transformer pre-processing
        if "tr_ln" in st:
            dxpre = _ln_bwd(plan, st["tr_ln"], g, None, A, "tr_ln_")
        elif "tr_drop" in st:
            dxpre = A("tr_drop_d", (M, D), dt)
            ops.epilogue_bwd(g, M, D, dout=dxpre, drop_p=st["tr_drop"]["p"], seed=st["tr_drop"]["seed"])
        else:
            dxpre = g
        # ---- positional conv: xpre = h0 + GELU(conv(h0) + bias)
        pcr = st["pc"]
        pc = pcr["pc"]
        _, K, Gn, Cg, Cgp = _posconv_geometry(pc, dt)
        Wd = Gn * Cgp
        front, pad, rows = pcr["front"], pcr["pad"], pcr["rows"]
        dz = A("pos_dz", (M, D), dt)
        ops.epilogue_bwd(dxpre, M, D, z=pcr["zpc"], dout=dz, dbias=_gb(plan, pc.conv.bias), act=1)
        dzp = A("pos_dzp", (rows, Wd), dt, zero=True)
        ops.posconv_pad(dz, dzp, B, T, Gn, Cg, Cgp, front, pad)
        norm, _, wdp, _ = plan.pos_w
        dh0 = A("dh0", (M, D), dt)
        off = front - (K - 1 - K // 2)
        if ops.posconv_tc_ok(dt, Cg, Cgp, K):
            ops.posconv_tc(dzp, wdp, None, dxpre, dh0, None, B, T, pc.embed_dim, Gn, Cg, Cgp, K, T + pad, off, 0)
        else:
            a = View(dzp, T, K * Cgp, Wd, kin=Cgp, kxs=Wd, sb0=Cgp, sb1=(T + pad) * Wd, offset=off * Wd)
            bv = View(wdp, Cg, K * Cgp, K * Cgp, kin=Cgp, kxs=Cgp, sb0=Cg * K * Cgp)
            G.gemm(a, bv, dh0, T, Cg, K * Cgp, nb0=Gn, nb1=B, ldc=D, c_sb0=Cg, c_sb1=T * D, residual=dxpre, ldr=D,
                   r_sb0=Cg, r_sb1=T * D)
        if pc.conv.weight_v.requires_grad:
            Kext = B * (T + pad)
            dwp = A("pos_dwp", (Gn, Cgp, K, Cgp), torch.float32)
            a = View(dzp, Cgp, Kext, 1, cs=Wd, sb0=Cgp, offset=front * Wd)
            bv = View(pcr["xp"], Cgp, Kext, 1, cs=Wd, sb0=Cgp, sb1=Wd)
            if dt == torch.bfloat16 and Cgp % 64 == 0:
                # all K taps folded into N (column = tap * Cgp + ci): one A tile feeds 4 taps of a 256-wide MMA instead
                # of K separate 64 x 64 products per group (which were bound by L2 -> shared-memory traffic)
                G.gemm(a, bv, dwp, Cgp, K * Cgp, Kext, nb0=Gn, ldc=K * Cgp, c_sb0=Cgp * K * Cgp, b_taps=K, cta_group=1,
                       block_n=256)
            else:
                G.gemm(a, bv, dwp, Cgp, Cgp, Kext, nb0=Gn, nb1=K, ldc=K * Cgp, c_sb0=Cgp * K * Cgp, c_sb1=Cgp)
            scratch = A("pos_wb_scratch", (K,), torch.float32)
            ops.posconv_weight_bwd(dwp, plan.view(pc.conv.weight_v), plan.view(pc.conv.weight_g), norm, scratch,
                                   plan.gview(pc.conv.weight_v), plan.gview(pc.conv.weight_g), D, Gn, Cg, Cgp, Cgp, K)
        dh0 = _acc(dh0, d_h0_head, M * D)
    else:
        dh0 = d_h0_head
    if dh0 is None:
        return None
    # ---- feature projection
    if st["lengths"] is not None:
        ops.pad_rows_zero(dh0, st["lengths"], B, T, D)
    fpr = st["fp"]
    gp = _out_proj_bwd(plan, dh0, None, None, None, fp.projection.bias, fpr["p_in"], fpr["seed_in"], dmasks, A, "fp_", M, D)
    dxn = A("fp_dxn", (M, C), dt)
    G.gemm(G.mat(gp), G.mat_t(plan.W(fp.projection.weight, dt)), dxn, M, C, D)
    _wgrad(plan, gp, fpr["xn"], fp.projection.weight, D, C, M)
    dfeat = _ln_bwd(plan, fpr["ln"], dxn, None, A, "fp_ln_")
    return dfeat.view(B, T, C)


# ----------------------------------------------------------------------------------------------------------------
# Whole-model forward / backward (no autograd inside: plain sequences of C-ABI launches)
# ----------------------------------------------------------------------------------------------------------------
def _out_lengths(model, lengths):
    """Integer frame arithmetic of ConvLayerBlock.forward (components.py:116-119)."""
    if lengths is None:
        return None
    blks = list(model.feature_extractor.conv_layers)
    lengths = lengths.contiguous()
    return ops.conv_out_lengths(lengths, torch.empty_like(lengths), [int(b.kernel_size) for b in blks],
                                [int(b.stride) for b in blks])


def forward_core(plan: Plan, wave, lengths, masks, A, train, save=True, num_layers=None, final_norm=False,
                 layer_keep=None):
    """Returns (hiddens [(B,T,D)], frame lengths, state). `masks`: {id(HardConcrete): fp32 mask tensor}."""
    model = plan.model
    _require_cuda(wave, "waveforms")
    if wave.dim() != 2:
        raise ValueError("Expected the input Tensor to be 2D (batch, time), but received %s" % list(wave.shape))
    dt = model.compute_dtype
    plan.refresh(dt)
    B, T0 = wave.shape
    if wave.dtype != torch.float32:
        raise _lib.DphError("waveforms must be float32")
    if lengths is not None:
        lengths = lengths.to(device=wave.device, dtype=torch.int64)
    if model.normalize_waveform:
        wn = A("wave_norm", (B, T0), torch.float32)
        ops.wave_norm(wave.contiguous(), wn, lengths, B, T0)
        wave = wn
    feat, st_conv = conv_stack_fwd(plan, wave, masks, save, A)
    flen = _out_lengths(model, lengths)
    if flen is not None:
        flen = flen.contiguous()
    hiddens, st_enc = encoder_fwd(plan, feat, flen, masks, A, train, num_layers=num_layers, final_norm=final_norm,
                                  layer_keep=layer_keep, save=save)
    return hiddens, flen, {"conv": st_conv, "enc": st_enc}


def backward_core(plan: Plan, st, dh, dmasks, A):
    """dh: per-hidden gradients ((M, D) / (B, T, D) compute dtype or None). Parameter gradients are accumulated into
    the gradient arena, HardConcrete mask gradients into `dmasks`."""
    plan.prepare_backward()
    enc = st["enc"]
    M = enc["B"] * enc["T"]
    dh = [None if g is None else g.view(M, -1) for g in dh]
    dfeat = encoder_bwd(plan, enc, dh, dmasks, A)
    if plan.grad_hook is not None:
        plan.grad_hook("encoder")
    if dfeat is not None:
        conv_stack_bwd(plan, st["conv"], dfeat, dmasks, A)
    join_side(plan)


class _ModelFn(torch.autograd.Function):
    """Autograd bridge for the module API: inputs are the mask vector (so that gradients reach `log_alpha` through the
    sampling node) — parameter gradients are written straight into `p.grad` (views of the gradient arena)."""

    @staticmethod
    def forward(ctx, plan, wave, lengths, mask_flat, masks, opts, anchor):
        A = FreshAlloc(plan.device)
        train = plan.model.training
        hiddens, flen, st = forward_core(plan, wave, lengths, masks, A, train, save=True, **opts)
        ctx.plan, ctx.st = plan, st
        ctx.has_masks = mask_flat is not None
        ctx.mark_non_differentiable(*([flen] if flen is not None else []))
        return tuple(hiddens) + ((flen,) if flen is not None else ())

    @staticmethod
    def backward(ctx, *grads):
        plan, st = ctx.plan, ctx.st
        n = st["enc"]["n_hidden"]
        dt = plan.model.compute_dtype
        dh = [None if g is None else g.to(dt).contiguous().clone() for g in grads[:n]]
        A = FreshAlloc(plan.device)
        dmasks = ops.zeros((max(plan.la_n, 1),), torch.float32, plan.device)
        backward_core(plan, st, dh, dmasks, A)
        ctx.st = None
        return None, None, None, (dmasks if ctx.has_masks else None), None, None, None


def _anchor(plan):
    """a differentiable scalar so autograd always runs our backward when any parameter requires grad"""
    for p in plan.params:
        if p.requires_grad:
            return p
    return None


def model_extract_features(model, waveforms, lengths=None, num_layers=None, _final=False):
    _require_cuda(waveforms, "waveforms")
    plan = get_plan(model)
    grad_on = torch.is_grad_enabled() and plan.any_grad
    tb = hc_tables(plan)
    opts = {"num_layers": num_layers, "final_norm": _final}
    if _final:
        tr = model.encoder.transformer
        if model.training and tr.layer_drop > 0:
            # LayerDrop: same host-side draw as the reference (components.py:902)
            opts["layer_keep"] = [not (torch.rand(1).item() <= tr.layer_drop) for _ in tr.layers]
    if not grad_on:
        masks = sample_masks(plan) if tb.nseg else {}
        masks = {k: v.detach() for k, v in masks.items()}
        A = FreshAlloc(plan.device)
        with torch.no_grad():
            hiddens, flen, _ = forward_core(plan, waveforms, lengths, masks, A, model.training, save=False, **opts)
        return list(hiddens), flen
    mask_flat = None
    masks = {}
    if tb.nseg:
        masks = sample_masks(plan)
        mask_flat = _flat_masks(plan, masks)
        masks = {k: v.detach() for k, v in masks.items()}
    outs = _ModelFn.apply(plan, waveforms, lengths, mask_flat, masks, opts, _anchor(plan))
    n = len(outs) - (1 if lengths is not None else 0)
    return list(outs[:n]), (outs[n] if lengths is not None else None)


def _flat_masks(plan, masks):
    """the (la_n,) tensor all mask views were sliced from (training: the sampling node's output)"""
    tb = hc_tables(plan)
    if plan.model.training:
        any_view = next(iter(masks.values()))
        base = any_view._base if any_view._base is not None else any_view
        return base
    flat = torch.zeros(plan.la_n, dtype=torch.float32, device=plan.device)
    return flat


def model_forward(model, waveforms, lengths=None):
    hiddens, flen = model_extract_features(model, waveforms, lengths, None, _final=True)
    x = hiddens[-1]
    if model.aux is not None:
        B, T, D = x.shape
        plan = get_plan(model)
        model = plan.model  # the padded shadow of a ragged pruned model, else the model itself
        out = torch.empty(B * T, model.aux.out_features, dtype=x.dtype, device=x.device)
        xd = x.detach().contiguous().view(B * T, D)
        G.gemm(G.mat(xd), G.mat(plan.W(model.aux.weight, x.dtype)), out, B * T, model.aux.out_features, D,
               bias=plan.view(model.aux.bias))
        x = out.view(B, T, -1)
    return x, flen
