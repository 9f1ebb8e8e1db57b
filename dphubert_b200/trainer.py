"""Native joint distill-and-prune training step (reference lightning.py:245-296 + optimizer lightning.py:200-238 +
gradient clipping distill.py:48 + DDP gradient all-reduce distill.py:41), as one static sequence of C-ABI launches:

    teacher forward (no grad)  ->  HardConcrete sampling  ->  student forward  ->  distillation heads + fused loss
    ->  expected size + Lagrangian  ->  backward (heads, encoder, conv stack, gates)  ->  gradient all-reduce (NCCL)
    ->  global-norm clip + AdamW over the flat parameter arenas (3 groups, lambda group with negative lr)

All buffers come from static workspaces, no step allocates, synchronises or reads back to the host, so the whole step
can be replayed from a CUDA graph (`capture=True`).  Step-varying scalars (learning rates, target sparsity, RNG epoch)
live in device memory and are refreshed outside the graph.
"""
import math
import os
from typing import List, Optional

import torch

from . import _lib, engine, ops
from . import gemm as G
from .engine import StaticAlloc, get_plan, hc_tables

ALIGN = engine.ALIGN


class FlatParams:
    """Small flat fp32 arena (+ gradient arena, + bf16 copy) for parameters living outside a model: the distillation
    heads and the two Lagrange multipliers."""

    def __init__(self, params: List[torch.nn.Parameter], device, align=ALIGN, garena=None):
        """`garena`: external gradient storage (a flat fp32 tensor of at least the arena's size) — the Lagrange
        multipliers keep their gradients in the tail of the student's gradient arena so that they ride in the same
        all-reduce call as the last student slice."""
        uniq = []
        seen = set()
        for p in params:
            if id(p) not in seen:
                seen.add(id(p))
                uniq.append(p)
        self.params = uniq
        self.offsets = {}
        off = 0
        for p in uniq:
            self.offsets[id(p)] = off
            off += (p.numel() + align - 1) // align * align
        self.total = (max(off, 1) + ALIGN - 1) // ALIGN * ALIGN
        self.device = device
        self.arena = ops.zeros((self.total,), torch.float32, device)
        if garena is not None:
            if garena.numel() < self.total:
                raise _lib.DphError("external gradient storage too small")
            self.garena = ops.memset(garena[: self.total])
        else:
            self.garena = ops.zeros((self.total,), torch.float32, device)
        self.a16 = torch.empty(self.total, dtype=torch.bfloat16, device=device)
        with torch.no_grad():
            for p in uniq:
                v = self.view(p)
                v.copy_(p.data.to(device))
                p.data = v
                p.grad = self.gview(p)
        self._vkey = None
        self.refresh()

    def view(self, p):
        o = self.offsets[id(p)]
        return self.arena[o:o + p.numel()].view(p.shape)

    def gview(self, p):
        o = self.offsets[id(p)]
        return self.garena[o:o + p.numel()].view(p.shape)

    def W(self, p, dtype):
        if dtype == torch.bfloat16:
            o = self.offsets[id(p)]
            return self.a16[o:o + p.numel()].view(p.shape)
        return self.view(p)

    def version(self):
        return sum(p._version for p in self.params)

    def refresh(self):
        ops.cast(self.arena, self.a16)
        self._vkey = self.version()

    def refresh_if_changed(self):
        """Re-cast the bf16 operand copy when a parameter was modified behind our back (load_state_dict, an in-place
        edit) — host-side check, so it has to run outside a captured graph."""
        if self._vkey != self.version():
            self.refresh()


def _head_linear(proj):
    """(Linear, has_gelu) of a distillation head: nn.Linear (layer2layer) or Sequential(Linear, GELU) (predlayer)."""
    if isinstance(proj, torch.nn.Linear):
        return proj, False
    if isinstance(proj, torch.nn.Sequential) and len(proj) == 2 and isinstance(proj[0], torch.nn.Linear) and \
            isinstance(proj[1], torch.nn.GELU):
        return proj[0], True
    raise ValueError("unsupported distillation head %r" % (proj,))


class NativeDistillStep:
    def __init__(self, teacher, student, distill_layers, distill_linear_projs, distill_mode="layer2layer",
                 l2_weight=0.0, l1_weight=1.0, cos_weight=1.0, cos_type="raw", use_reg=True, lambdas=None,
                 original_num_params=None, learning_rate=2e-4, weight_decay=0.0, reg_learning_rate=0.02,
                 warmup_updates=15000, max_updates=50000, target_sparsity=0.75, sparsity_warmup_updates=5000,
                 clip_norm=10.0, betas=(0.9, 0.999), eps=1e-8, process_group=None):
        if distill_mode not in ("layer2layer", "predlayer"):
            raise ValueError(f"Invalid distill mode: {distill_mode}")
        assert len(distill_layers) == len(distill_linear_projs)
        assert cos_type in ("raw", "log_sig"), cos_type
        self.teacher, self.student = teacher, student
        self.layers = list(distill_layers)
        self.projs = list(distill_linear_projs)
        self.mode = distill_mode
        self.w2, self.w1, self.wc = float(l2_weight), float(l1_weight), float(cos_weight)
        self.cos_type = 0 if cos_type == "raw" else 1
        self.use_reg = bool(use_reg)
        self.lr, self.wd, self.reg_lr = float(learning_rate), float(weight_decay), float(reg_learning_rate or 0.0)
        self.warmup, self.max_updates = int(warmup_updates), int(max_updates)
        self.target_sparsity = float(target_sparsity or 0.0)
        self.sparsity_warmup = int(sparsity_warmup_updates or 0)
        self.clip_norm = float(clip_norm or 0.0)
        self.betas, self.eps = betas, eps
        self.pg = process_group
        self.world = torch.distributed.get_world_size(process_group) if (
            torch.distributed.is_available() and torch.distributed.is_initialized()) else 1

        self.tplan = get_plan(teacher)
        self.splan = get_plan(student)
        self._pending, self._covered = [], set()
        self.splan.grad_hook = self._grads_ready if self.world > 1 else None
        self.two_streams = True   # teacher forward / weight-gradient GEMMs on side streams (CUDA only)
        # transformer layers per gradient all-reduce (DPH_AR_GROUP; measured at 2 GPUs: 1 -> +0.54 ms over one GPU, 3 -> +0.42)
        self.ar_group = max(1, int(os.environ.get("DPH_AR_GROUP", "3")))
        self.tstream = None
        dev = self.splan.device
        self.device = dev
        if teacher.compute_dtype != student.compute_dtype:
            raise _lib.DphError("teacher and student must use the same compute dtype")
        self.dt = student.compute_dtype
        head_params = []
        for pr in self.projs:
            lin, _ = _head_linear(pr)
            head_params += [lin.weight, lin.bias]
        if self.use_reg:
            if lambdas is None:
                lambdas = [torch.nn.Parameter(torch.tensor(0.0)), torch.nn.Parameter(torch.tensor(0.0))]
            self.lambdas = list(lambdas)
        else:
            self.lambdas = []
        self.heads = FlatParams(head_params, dev)
        self.symm = None   # symmetric-memory handle of the student's gradient arena (multicast all-reduce), else NCCL
        # DPH_ALLREDUCE = multimem | nccl | auto (default): the switch-side reduction when the box offers a multicast
        # address, else NCCL.  Measured on 8 B200: 9.19 ms per step against 9.66 with NCCL (one GPU: 8.90).
        # On 2 GPUs NCCL's direct peer copies are the faster exchange (9.30 against 9.42 ms), from 4 GPUs on the switch
        # wins (4: 9.26 against 9.51; 8: 9.19 against 9.66): `auto` picks accordingly.
        mode = os.environ.get("DPH_ALLREDUCE", "auto")
        if self.world > 1 and dev.type == "cuda" and (mode == "multimem" or (mode == "auto" and self.world >= 3)):
            try:
                self._setup_multimem()
            except Exception as e:  # noqa: BLE001  (every rank sees the same system, so every rank lands here together)
                if mode == "multimem":
                    raise
                import sys
                print("dphubert_b200: multicast all-reduce unavailable (%s); using NCCL" % (e,), file=sys.stderr)
                self.symm = None
        self.lam = FlatParams(self.lambdas, dev, align=1, garena=self.splan.gtail) if self.use_reg else None
        if original_num_params is None:
            original_num_params = sum(p.numel() for p in teacher.parameters())
        self.n0 = float(original_num_params)

        self.At = StaticAlloc(dev, "t_")
        self.As = StaticAlloc(dev, "s_")
        self.Ah = StaticAlloc(dev, "h_")
        self.global_step = 0
        # device-side step-varying scalars
        self.lr_dev = ops.zeros((4,), torch.float32, dev)       # [main, log_alpha, lambda, target sparsity]
        # ring of pinned staging slots, each guarded by an event: the host runs ahead of the device (replay() never
        # synchronises), so a single slot could be overwritten before its queued H2D copy has executed
        cuda = torch.cuda.is_available() and self.device.type == "cuda"
        self._ring = [(torch.zeros(4, dtype=torch.float32).pin_memory() if cuda else torch.zeros(4, dtype=torch.float32),
                       torch.cuda.Event() if cuda else None) for _ in range(16)]
        self._ring_used = [False] * 16
        self._ring_i = 0
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.gnorm = torch.zeros(1, dtype=torch.float64, device=dev)
        self.out = ops.zeros((16,), torch.float32, dev)  # [0:4] distill loss parts, [8:13] lagrangian block
        self._opt_state = None
        # gate-noise stream: seeded from the host seed stream (ops.manual_seed / next_seed), saved in state_dict()
        self._hc_seed = ops.next_seed()
        self._graph = None
        self._gcache, self._gcache_gen, self._shape_seen = {}, None, {}
        self.graph_capacity = 128   # graphs kept by `step()` (least recently used dropped first); 0 disables the cache
        self._static_in = None
        self.comm_stream = None

    # ------------------------------------------------------------------------------------------------ schedules
    def _lr_scale(self, step_count):
        """LinearDecayLRScheduler.get_lr (lightning.py:37-44) as a scale factor; `step_count` = scheduler._step_count
        for the update about to be applied (1 on the first optimizer step)."""
        if step_count <= self.warmup:
            return step_count / self.warmup if self.warmup > 0 else 1.0
        if step_count >= self.max_updates:
            return 0.0
        return (self.max_updates - step_count) / (self.max_updates - self.warmup)

    def _target(self):
        """DistillModule._get_target_sparsity (lightning.py:240-243)."""
        if self.global_step >= self.sparsity_warmup:
            return self.target_sparsity
        return self.target_sparsity * (self.global_step / self.sparsity_warmup)

    def _push_scalars(self):
        sc = self._lr_scale(self.global_step + 1)
        i = self._ring_i % len(self._ring)
        self._ring_i += 1
        host, ev = self._ring[i]
        if ev is not None and self._ring_used[i]:
            ev.synchronize()   # the copy queued from this slot 16 steps ago has executed
        host[0] = self.lr * sc
        host[1] = self.reg_lr * sc
        host[2] = -self.reg_lr * sc
        host[3] = self._target() if self.use_reg else 0.0
        self.lr_dev.copy_(host, non_blocking=True)
        if ev is not None:
            ev.record()
            self._ring_used[i] = True
        # host-side staleness checks that a captured graph cannot replay: derived GEMM operands follow
        # parameters changed by load_state_dict / in-place edits
        self.heads.refresh_if_changed() if self.dt == torch.bfloat16 else None
        if self._graph is not None or self._gcache:
            self.splan.refresh(self.dt)
            self.tplan.refresh(self.dt)

    # ------------------------------------------------------------------------------------------------ the step body
    def _sample_masks(self):
        """All HardConcrete gates of the student in one launch (hardconcrete.py:92-99 x 31 modules)."""
        plan = self.splan
        tb = hc_tables(plan)
        if tb.nseg == 0:
            return {}, None
        la = plan.arena[plan.la_off:]
        masks = self.As("hc_masks", (plan.la_n,), torch.float32)
        u = self.As("hc_u", (plan.la_n,), torch.float32)
        injected = any(h.noise is not None for h in tb.hcs)
        if injected:
            # parity hook: identical uniform noise as the oracle (every gate must have been given noise)
            for h in tb.hcs:
                o = plan.offsets[id(h.log_alpha)] - plan.la_off
                u[o:o + h.n_in].copy_(h.noise.to(plan.device, torch.float32))
                h.noise = None
            ops.hardconcrete_fwd(la, u, None, masks, plan.la_n, tb.beta, tb.l, tb.r, tb.eps, 0)
        else:
            ops.hardconcrete_fwd(la, None, u, masks, plan.la_n, tb.beta, tb.l, tb.r, tb.eps, self._hc_seed)
        out = {}
        for h in tb.hcs:
            o = plan.offsets[id(h.log_alpha)] - plan.la_off
            out[id(h)] = masks[o:o + h.n_in]
        return out, u

    def _expected_size(self):
        """Expected size + Lagrangian (lightning.py:267-271): (gate tables, regulariser on?, d n_hat / d norms)."""
        splan = self.splan
        tb = hc_tables(splan)
        reg_on = self.use_reg and tb.nseg > 0
        dnorm = None
        if reg_on:
            A = self.Ah
            la = splan.arena[splan.la_off:]
            norms = A("l0_norms", (tb.nseg,), torch.float32)
            nhat = A("n_hat", (1,), torch.float32)
            dnorm = A("dnorm", (tb.nseg,), torch.float32)
            ops.l0_norms(la, tb.seg_off, tb.seg_len, tb.nseg, tb.bias, norms)
            ops.num_params(tb.prog, norms, tb.nseg, nhat, dnorm)
            ops.lagrangian(nhat, self.lam.arena, self.n0, self.lr_dev[3:4], self.out[8:13])
        return tb, reg_on, dnorm

    def _body(self, wave, lengths):
        dt = self.dt
        splan, tplan = self.splan, self.tplan
        ops.rng_advance()
        # ---- teacher (eval, no grad) on its own stream, concurrently with the student forward: the two passes share
        # nothing but the input waveform, and each one's kernel tails / small kernels fill the other's bubbles
        if wave.is_cuda and self.two_streams:
            if self.tstream is None:
                self.tstream = torch.cuda.Stream()
                splan.side_stream = torch.cuda.Stream()
            main = torch.cuda.current_stream()
            self.tstream.wait_stream(main)
            with torch.cuda.stream(self.tstream):
                # the regulariser depends on log_alpha only: its three small, latency-bound kernels run here, off the
                # student's critical path (they used to sit between the loss forward and the loss backward)
                reg = self._expected_size()
                t_h, _, _ = engine.forward_core(tplan, wave, lengths, {}, self.At, False, save=False)
        else:
            main = None
            reg = self._expected_size()
            t_h, _, _ = engine.forward_core(tplan, wave, lengths, {}, self.At, False, save=False)
        # ---- student
        masks, u = self._sample_masks()
        s_h, _, st = engine.forward_core(splan, wave, lengths, masks, self.As, self.student.training, save=True)
        if main is not None:
            main.wait_stream(self.tstream)
        B, T, D = s_h[0].shape
        M = B * T
        nL = len(self.layers)
        total_rows = M * nL
        # ---- distillation heads + fused loss (lightning.py:255-265)
        A = self.Ah
        acc = A("loss_acc", (4,), torch.float32)
        preds, recs = [], []
        for i, (idx, pr) in enumerate(zip(self.layers, self.projs)):
            lin, gelu = _head_linear(pr)
            src = s_h[idx] if self.mode == "layer2layer" else s_h[-1]
            src_i = idx if self.mode == "layer2layer" else len(s_h) - 1
            Dt = lin.out_features
            pred = A("pred%d" % i, (M, Dt), dt)
            pre = A("pred_pre%d" % i, (M, Dt), dt) if gelu else None
            G.gemm(G.mat(src.view(M, D)), G.mat(self.heads.W(lin.weight, dt)), pred, M, Dt, D,
                   bias=self.heads.view(lin.bias), act=1 if gelu else 0, c2=pre)
            rowstat = A("rowstat%d" % i, (M * 3,), torch.float32)
            tgt = t_h[idx].view(M, Dt)
            ops.distill_loss_fwd(pred, tgt, rowstat, acc, self.out, M, Dt, self.w2, self.w1, self.wc, self.cos_type,
                                 i == 0, i == nL - 1, total_rows)
            preds.append(pred)
            recs.append((lin, gelu, pre, rowstat, tgt, src_i, Dt))
        tb, reg_on, dnorm = reg
        # ---- backward: heads
        splan.prepare_backward()
        dh = [None] * len(s_h)
        for i in range(nL - 1, -1, -1):
            lin, gelu, pre, rowstat, tgt, src_i, Dt = recs[i]
            dpred = A("dpred%d" % i, (M, Dt), dt)
            ops.distill_loss_bwd(preds[i], tgt, rowstat, None, dpred, M, total_rows, Dt, self.w2, self.w1, self.wc,
                                 self.cos_type)
            gb = self.heads.gview(lin.bias)
            if gelu:
                ops.epilogue_bwd(dpred, M, Dt, z=pre, dout=dpred, dbias=gb, act=1)
            else:
                ops.epilogue_bwd(dpred, M, Dt, dbias=gb)
            src = s_h[src_i].view(M, D)
            with engine._on_side(splan):
                G.gemm(G.mat_t(dpred), G.mat_t(src), self.heads.gview(lin.weight), Dt, D, M, atomic=True,
                       ksplit=engine._ksplit(Dt, D, M))
            Wt = G.mat_t(self.heads.W(lin.weight, dt))
            if dh[src_i] is None:
                dh[src_i] = A("dhid%d" % src_i, (M, D), dt)
                G.gemm(G.mat(dpred), Wt, dh[src_i], M, D, Dt)
            else:
                G.gemm(G.mat(dpred), Wt, dh[src_i], M, D, Dt, accumulate=True)
        self._heads_ready()   # the heads' gradients are complete: their all-reduce overlaps the whole student backward
        # ---- backward: student
        dmasks = self.As("hc_dmasks", (max(splan.la_n, 1),), torch.float32)
        ops.memset(dmasks)
        engine.backward_core(splan, st, dh, dmasks, self.As)
        if tb.nseg:
            la = splan.arena[splan.la_off:]
            dla = splan.garena[splan.la_off:]
            ops.hardconcrete_bwd(la, u, dmasks, dla, splan.la_n, tb.beta, tb.l, tb.r)
            if reg_on:
                ops.l0_norms_bwd(la, dla, tb.seg_off, tb.seg_len, tb.nseg, tb.bias, dnorm, None, self.out[10:11])
                # d loss / d lambda1 = (s - t), d loss / d lambda2 = (s - t)^2
                ops.axpy(self.lam.garena[0:1], self.out[11:12], 1)
                ops.axpy(self.lam.garena[1:2], self.out[12:13], 1)

    def _grad_buffers(self):
        """the distinct gradient storages: student arena + its tail (which holds the multipliers' gradients), heads"""
        return [self.splan.garena_full, self.heads.garena]

    def arenas(self):
        """[(param arena, grad arena)] of everything trainable."""
        out = [(self.splan.arena, self.splan.garena), (self.heads.arena, self.heads.garena)]
        if self.lam is not None:
            out.append((self.lam.arena, self.lam.garena))
        return out

    # ------------------------------------------------------------------------------------------------ optimizer
    def _untouched(self):
        """ids of student parameters the reference's optimizer never updates because their `.grad` stays None
        (torch.optim.AdamW skips them, so they must not be weight-decayed either): frozen tensors such as
        `feature_extractor.dummy_weight`, the Transformer-level LayerNorm of pre-norm models (unused by
        `extract_features`, components.py:885-892), the `aux` head, and — in layer2layer mode — the transformer layers
        above the last distilled one."""
        st = self.student
        tr = st.encoder.transformer
        skip = {id(p) for p in st.parameters() if not p.requires_grad}
        if not tr.layer_norm_first:
            skip |= {id(p) for p in tr.layer_norm.parameters()}
        if st.aux is not None:
            skip |= {id(p) for p in st.aux.parameters()}
        if self.mode == "layer2layer":
            for li, layer in enumerate(tr.layers):
                if li >= max(self.layers):
                    skip |= {id(p) for p in layer.parameters()}
        return skip

    def _runs(self, lo, hi, skip):
        """maximal [a, b) runs of the student arena range [lo, hi) made of parameters the optimizer updates"""
        plan = self.splan
        runs = []
        for p in plan.params:
            o = plan.offsets[id(p)]
            if o < lo or o >= hi or id(p) in skip:
                continue
            e = o + (p.numel() + ALIGN - 1) // ALIGN * ALIGN
            if runs and runs[-1][1] == o:
                runs[-1][1] = e
            else:
                runs.append([o, e])
        return runs

    def _segments(self):
        """(arena key, start, end, lr index, weight decay, has bf16 mirror) per optimizer launch.  With weight decay
        off (run.sh default) a parameter with a zero gradient is left bit-identical by AdamW, so whole regions go in
        one launch; with weight decay on, the regions are cut into runs of parameters the reference really updates."""
        sp = self.splan
        bf = self.dt == torch.bfloat16
        regions = [(0, sp.lin_total, 0, self.wd, bf), (sp.lin_total, sp.la_off, 0, self.wd, False),
                   (sp.la_off, sp.total, 1, 0.0, False)]
        segs = []
        skip = self._untouched() if self.wd != 0.0 else None
        for lo, hi, li, wd, mirror in regions:
            if hi <= lo:
                continue
            runs = [[lo, hi]] if (skip is None or wd == 0.0) else self._runs(lo, hi, skip)
            for a, b in runs:
                segs.append(("student", a, b, li, wd, mirror))
        segs.append(("heads", 0, self.heads.total, 0, self.wd, bf))
        if self.lam is not None:
            segs.append(("lam", 0, 2, 2, 0.0, False))
        return segs

    def _arena_of(self, key):
        if key == "student":
            return self.splan.arena, self.splan.garena, self.splan.lin16
        if key == "heads":
            return self.heads.arena, self.heads.garena, self.heads.a16
        return self.lam.arena, self.lam.garena, None

    def _ensure_opt_state(self):
        if self._opt_state is None:
            keys = ["student", "heads"] + (["lam"] if self.lam is not None else [])
            self._opt_state = {k: (ops.zeros(self._arena_of(k)[0].shape, torch.float32, self.device),
                                   ops.zeros(self._arena_of(k)[0].shape, torch.float32, self.device)) for k in keys}
            self._segs = self._segments()
        return self._opt_state

    def _optimizer_step(self):
        """clip_grad_norm_(max_norm) + AdamW over the arenas (lightning.py:200-238, distill.py:48)."""
        state = self._ensure_opt_state()
        segs = self._segs
        ops.counter_inc(self.step_dev)
        use_clip = self.clip_norm > 0
        if use_clip:
            # the norm covers every gradient (untouched parameters have zero gradients)
            ops.memset(self.gnorm)
            for g in self._grad_buffers():
                ops.sumsq(g, self.gnorm)
        for key, a, b, li, wd, mirror in segs:
            p, g, p16 = self._arena_of(key)
            m, v = state[key]
            ops.adamw(p[a:b], g[a:b], m[a:b], v[a:b], self.lr_dev[li:li + 1], self.betas[0], self.betas[1], self.eps, wd,
                      self.step_dev, self.gnorm if use_clip else None, self.clip_norm,
                      p16[a:b] if (mirror and p16 is not None) else None, True)
        if self.wd != 0.0:
            # gradients of skipped runs were not cleared by the optimizer launches
            for _, g in self.arenas():
                ops.memset(g)
        # derived GEMM operands that are not plain casts (tap-major conv weights, weight-normalised pos conv)
        self.splan.rebuild_derived(self.dt, cast_lin=False)

    # ------------------------------------------------------------------------------------------------ checkpoint state
    def _named_slots(self):
        """[(name, arena key, offset, numel)] of every parameter the optimizer owns, under the reference's
        Lightning-level names (`student_model.`, `distill_linear_projs.N.`, `lambda1/2`)."""
        out = []
        for n, p in self.student.named_parameters():
            out.append(("student_model." + n, "student", self.splan.offsets[id(p)], p.numel()))
        seen = set()
        for i, pr in enumerate(self.projs):
            lin, gelu = _head_linear(pr)
            if id(lin) in seen:
                continue
            seen.add(id(lin))
            pre = "distill_linear_projs.%d.%s" % (i, "0." if gelu else "")
            out.append((pre + "weight", "heads", self.heads.offsets[id(lin.weight)], lin.weight.numel()))
            out.append((pre + "bias", "heads", self.heads.offsets[id(lin.bias)], lin.bias.numel()))
        for i, l in enumerate(self.lambdas):
            out.append(("lambda%d" % (i + 1), "lam", self.lam.offsets[id(l)], 1))
        return out

    def state_dict(self):
        """Everything `trainer.fit(ckpt_path=...)` restores in the reference besides the weights (distill.py:141-144,
        180-185): Adam moments per parameter (by name), the optimizer / scheduler step counters, `global_step`, and
        the state of the noise streams (gate-noise seed, host seed counter, device RNG epoch).  Synchronises."""
        state = self._ensure_opt_state()
        moments = {}
        for name, key, off, n in self._named_slots():
            m, v = state[key]
            moments[name] = {"exp_avg": m[off:off + n].detach().cpu().clone(),
                             "exp_avg_sq": v[off:off + n].detach().cpu().clone()}
        return {"global_step": int(self.global_step), "optimizer_step": int(self.step_dev.cpu()[0]),
                "moments": moments, "hc_seed": int(self._hc_seed), "seed_counter": ops.seed_state(),
                "rng_epoch": ops.rng_get(self.device), "format": 1}

    def load_state_dict(self, sd):
        """Inverse of `state_dict()`; weights are restored separately (module.load_state_dict).  A captured graph
        stays valid: every buffer is updated in place."""
        state = self._ensure_opt_state()
        slots = {name: (key, off, n) for name, key, off, n in self._named_slots()}
        missing = [k for k in slots if k not in sd["moments"]]
        extra = [k for k in sd["moments"] if k not in slots]
        if missing or extra:
            raise KeyError("optimizer state does not match the model: missing %s unexpected %s" % (missing[:5], extra[:5]))
        for name, (key, off, n) in slots.items():
            m, v = state[key]
            ent = sd["moments"][name]
            if ent["exp_avg"].numel() != n:
                raise ValueError("optimizer state of %s has %d elements, expected %d" % (name, ent["exp_avg"].numel(), n))
            m[off:off + n].copy_(ent["exp_avg"].flatten().to(self.device))
            v[off:off + n].copy_(ent["exp_avg_sq"].flatten().to(self.device))
        self.global_step = int(sd["global_step"])
        self.step_dev.fill_(int(sd["optimizer_step"]))
        self._hc_seed = int(sd["hc_seed"])
        ops.set_seed_state(sd["seed_counter"])
        ops.rng_set(int(sd["rng_epoch"]))
        if (self._graph is not None or self._gcache) and int(sd["hc_seed"]) != self._captured_hc_seed:
            raise _lib.DphError("the gate-noise seed is baked into the captured graph: load the state before capture()")
        # parameters were (re)loaded by the caller: rebuild every derived operand from them
        self.heads.refresh()
        self.splan.mark_dirty()
        self.splan.refresh(self.dt)

    # ------------------------------------------------------------------------------------------------ data parallel
    # DDP's gradient mean (distill.py:41 strategy="ddp"): the Linear-weight gradients of transformer layer l (one
    # contiguous 28 MB slice of the gradient arena for HuBERT Base) are all-reduced as soon as the backward pass has
    # left that layer, on NCCL's own stream, while the earlier layers' backward keeps the SMs busy; everything else
    # (conv stack, biases, norms, log_alpha, heads, lambdas) follows after the backward. Works inside CUDA graph capture.
    def _setup_multimem(self):
        """Put the student's gradient arena into symmetric memory bound to an NVSwitch multicast object
        (torch.distributed._symmetric_memory does the allocation and the handle exchange); ranks then reduce slices of
        it with `dph_multimem_allreduce` instead of NCCL.  Collective: every rank must call it."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        plan = self.splan
        n = plan.garena_full.numel()
        buf = symm_mem.empty(n, dtype=torch.float32, device=self.device)
        hdl = symm_mem.rendezvous(buf, self.pg if self.pg is not None else dist.group.WORLD)
        if not hdl.multicast_ptr:
            raise _lib.DphError("DPH_ALLREDUCE=multimem: this system gave the symmetric allocation no multicast address")
        plan.rebind_grad_arena(buf)
        self.symm = hdl
        self.mm_ctas = int(os.environ.get("DPH_MM_CTAS", "8"))

    def _reduce_multimem(self, t):
        """All-reduce (mean) of a slice of the symmetric gradient arena on the current stream: barrier (every rank's
        slice is complete), switch-side reduction of this rank's share, barrier (every mean has landed)."""
        full = self.splan.garena_full
        off = (t.data_ptr() - full.data_ptr()) // 4
        n = t.numel()
        lo, hi = off & ~3, (off + n + 3) & ~3          # 16-byte granularity (arena offsets are aligned anyway)
        hi = min(hi, full.numel() & ~3)
        self.symm.barrier(0)
        ops.multimem_allreduce(self.symm.multicast_ptr + 4 * lo, hi - lo, self.symm.rank, self.symm.world_size,
                               1.0 / self.world, self.mm_ctas)
        self.symm.barrier(0)

    def _reduce_async(self, t):
        import torch.distributed as dist

        if t.numel() == 0:
            return
        if self.symm is not None and t.untyped_storage().data_ptr() == self.splan.garena_full.untyped_storage().data_ptr():
            self._reduce_multimem(t)
            return
        if t.is_cuda:
            w = dist.all_reduce(t, op=dist.ReduceOp.AVG, group=self.pg, async_op=True)
            self._pending.append((w, None))
        else:  # gloo (CPU tests): no AVG
            w = dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
            self._pending.append((w, t))

    def _grads_ready(self, key):
        """Backward-pass hook (engine.Plan.grad_ranges): every gradient of slice `key` — a transformer layer,
        "encoder" (projection / positional conv / transformer-level vectors) or "conv_hi" (conv layers n-1 .. 1) — has
        been issued; reduce it now, behind both streams that produced it, while the rest of the backward runs."""
        rng = self.splan.grad_ranges.get(key)
        if rng is None or key in self._covered:
            return
        if isinstance(key, int) and self.ar_group > 1:
            # transformer layers complete from the top down and lie in that order in the arena: reduce them in groups of
            # `ar_group` layers (fewer, larger collectives), at the moment the lowest layer of a group is done
            if key % self.ar_group != 0:
                return
            keys = [k for k in range(key, key + self.ar_group) if self.splan.grad_ranges.get(k) is not None
                    and k not in self._covered]
            lo = min(self.splan.grad_ranges[k][0] for k in keys)
            hi = max(self.splan.grad_ranges[k][1] for k in keys)
            self._reduce_slice(self.splan.garena_full[lo:hi])
            self._covered.update(keys)
            return
        self._reduce_slice(self.splan.garena_full[rng[0]:rng[1]])
        self._covered.add(key)

    def _reduce_slice(self, t):
        side = self.splan.side_stream
        if side is not None and t.is_cuda:
            # weight gradients come from the side stream, bias / norm gradients from the main one
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._reduce_async(t)
        else:
            self._reduce_async(t)

    def _heads_ready(self):
        if self.world > 1:
            self._reduce_slice(self.heads.garena)

    def _allreduce(self):
        """The last slice (conv layer 0, every log_alpha and — in the gradient tail — the two multipliers: one call),
        any slice whose hook did not fire, then wait for everything."""
        if self.world <= 1:
            return
        plan = self.splan
        rest = [k for k in plan.grad_ranges if k != "tail" and k not in self._covered and plan.grad_ranges[k] is not None]
        for k in rest:
            self._reduce_async(plan.garena_full[plan.grad_ranges[k][0]:plan.grad_ranges[k][1]])
        lo, hi = plan.grad_ranges["tail"]
        self._reduce_async(plan.garena_full[lo:hi])
        for w, t in self._pending:
            w.wait()
            if t is not None:
                t.mul_(1.0 / self.world)
        self._pending, self._covered = [], set()

    # ------------------------------------------------------------------------------------------------ public
    def zero_grad(self):
        for _, g in self.arenas():
            ops.memset(g)

    def run_step(self, wave, lengths=None):
        """One full training step on device-resident inputs. Returns the (16,) device vector of loss terms:
        [0] distill total, [1] mse, [2] l1, [3] cos, [8] reg loss, [9] expected sparsity."""
        self._push_scalars()
        self._body(wave, lengths)
        self._allreduce()
        self._optimizer_step()
        self.global_step += 1
        return self.out

    # ------------------------------------------------------------------------------------------------ shape-keyed graphs
    def step(self, wave, lengths=None):
        """`run_step` for variable-shape batches without the host cost of an eager launch sequence.

        The reference crops every batch to its shortest utterance (audio_dataset.py:285-363), so (B, T) changes from
        batch to batch and one captured CUDA graph cannot serve; issued eagerly, the ~480 launches of a step cost more
        host time than the GPU needs to run them.  This entry point keeps a cache of captured graphs keyed by the batch
        shape: a shape runs eagerly the first time it is seen (which sizes the workspace), is captured the second time,
        and is replayed from then on.  All graphs share the one name-keyed workspace (their staging buffers are cleared
        inside each graph); if a later, larger shape makes a workspace buffer grow, the cached graphs hold stale
        addresses and are dropped.  Use `DeviceLoader(..., quantum=...)` so that shapes actually recur."""
        if not wave.is_cuda or self.graph_capacity <= 0:
            return self.run_step(wave, lengths)
        allocs = (self.At, self.As, self.Ah)

        def eager():
            # graph replays rewrite the shared staging buffers behind the allocator's back: an eager step in between
            # must clear them itself instead of trusting "same shape as my last request"
            for a in allocs:
                a.always_zero = True
            try:
                return self.run_step(wave, lengths)
            finally:
                for a in allocs:
                    a.always_zero = False

        key = (tuple(wave.shape), lengths is not None)
        gen = tuple(a.generation for a in allocs)
        if gen != self._gcache_gen:
            self._gcache.clear()
            self._gcache_gen = gen
        ent = self._gcache.get(key)
        if ent is None:
            seen = self._shape_seen.get(key, 0)
            self._shape_seen[key] = seen + 1
            if seen == 0 or self._opt_state is None:
                return eager()
            w_in = self.As("in_wave", tuple(wave.shape), torch.float32)
            l_in = self.As("in_lengths", (wave.shape[0],), torch.int64) if lengths is not None else None
            if tuple(a.generation for a in allocs) != gen:   # the input buffers themselves grew: next time
                return eager()
            for a in allocs:
                a.always_zero = True
            try:
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._body(w_in, l_in)
                    self._allreduce()
                    self._optimizer_step()
            finally:
                for a in allocs:
                    a.always_zero = False
            if tuple(a.generation for a in allocs) != gen:
                raise _lib.DphError("a workspace buffer was re-allocated during graph capture")
            self._captured_hc_seed = self._hc_seed
            ent = (g, w_in, l_in)
            self._gcache[key] = ent
            while len(self._gcache) > self.graph_capacity:
                self._gcache.pop(next(iter(self._gcache)))
        else:
            self._gcache.pop(key)      # re-insert: most recently used last
            self._gcache[key] = ent
        g, w_in, l_in = ent
        w_in.copy_(wave, non_blocking=True)
        if l_in is not None:
            l_in.copy_(lengths, non_blocking=True)
        self._push_scalars()
        g.replay()
        self.global_step += 1
        return self.out

    def capture(self, wave, lengths=None):
        """Capture the step (fixed shapes) into a CUDA graph. `wave`/`lengths` become the static input buffers."""
        self._static_in = (wave, lengths)
        # warm-up outside capture: allocates every static buffer and the optimizer state
        self._push_scalars()
        self._body(wave, lengths)
        self._allreduce()
        self._optimizer_step()
        self.global_step += 1
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._body(wave, lengths)
            self._allreduce()
            self._optimizer_step()
        self._graph = g
        self._captured_hc_seed = self._hc_seed
        for a in (self.At, self.As, self.Ah):
            a.frozen = True
        return g

    def replay(self):
        self._push_scalars()
        self._graph.replay()
        self.global_step += 1
        return self.out

    def loss_terms(self):
        """Host copy of the loss terms of the last step (synchronises)."""
        o = self.out.detach().cpu()
        reg = float(o[8]) if self.use_reg else 0.0
        return {"loss": float(o[0]) + reg, "loss_distill": float(o[0]), "loss_mse": float(o[1]), "loss_l1": float(o[2]),
                "loss_cos": float(o[3]), "loss_reg": reg, "sparsity_expected": float(o[9]) if self.use_reg else None}
