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
from typing import List, Optional

import torch

from . import _lib, engine, ops
from . import gemm as G
from .engine import StaticAlloc, get_plan, hc_tables

ALIGN = engine.ALIGN


class FlatParams:
    """Small flat fp32 arena (+ gradient arena, + bf16 copy) for parameters living outside a model: the distillation
    heads and the two Lagrange multipliers."""

    def __init__(self, params: List[torch.nn.Parameter], device, align=ALIGN):
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
        self.garena = ops.zeros((self.total,), torch.float32, device)
        self.a16 = torch.empty(self.total, dtype=torch.bfloat16, device=device)
        with torch.no_grad():
            for p in uniq:
                v = self.view(p)
                v.copy_(p.data.to(device))
                p.data = v
                p.grad = self.gview(p)
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

    def refresh(self):
        ops.cast(self.arena, self.a16)


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
        self._pending, self._covered = [], []
        self.splan.grad_hook = self._layer_grads_ready if self.world > 1 else None
        self.two_streams = True   # teacher forward / weight-gradient GEMMs on side streams (CUDA only)
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
        self.lam = FlatParams(self.lambdas, dev, align=1) if self.use_reg else None
        if original_num_params is None:
            original_num_params = sum(p.numel() for p in teacher.parameters())
        self.n0 = float(original_num_params)

        self.At = StaticAlloc(dev, "t_")
        self.As = StaticAlloc(dev, "s_")
        self.Ah = StaticAlloc(dev, "h_")
        self.global_step = 0
        # device-side step-varying scalars
        self.lr_dev = ops.zeros((4,), torch.float32, dev)       # [main, log_alpha, lambda, target sparsity]
        self.lr_host = torch.zeros(4, dtype=torch.float32).pin_memory() if torch.cuda.is_available() else \
            torch.zeros(4, dtype=torch.float32)
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.gnorm = torch.zeros(1, dtype=torch.float64, device=dev)
        self.out = ops.zeros((16,), torch.float32, dev)  # [0:4] distill loss parts, [8:13] lagrangian block
        self._opt_state = None
        self._graph = None
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
        self.lr_host[0] = self.lr * sc
        self.lr_host[1] = self.reg_lr * sc
        self.lr_host[2] = -self.reg_lr * sc
        self.lr_host[3] = self._target() if self.use_reg else 0.0
        self.lr_dev.copy_(self.lr_host, non_blocking=True)

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
                t_h, _, _ = engine.forward_core(tplan, wave, lengths, {}, self.At, False, save=False)
        else:
            main = None
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
        # ---- expected size + Lagrangian (lightning.py:267-271)
        tb = hc_tables(splan)
        reg_on = self.use_reg and tb.nseg > 0
        if reg_on:
            la = splan.arena[splan.la_off:]
            norms = A("l0_norms", (tb.nseg,), torch.float32)
            nhat = A("n_hat", (1,), torch.float32)
            dnorm = A("dnorm", (tb.nseg,), torch.float32)
            ops.l0_norms(la, tb.seg_off, tb.seg_len, tb.nseg, tb.bias, norms)
            ops.num_params(tb.prog, norms, tb.nseg, nhat, dnorm)
            ops.lagrangian(nhat, self.lam.arena, self.n0, self.lr_dev[3:4], self.out[8:13])
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

    @property
    def _hc_seed(self):
        return 0x1234ABCD

    def arenas(self):
        """[(param arena, grad arena)] of everything trainable."""
        out = [(self.splan.arena, self.splan.garena), (self.heads.arena, self.heads.garena)]
        if self.lam is not None:
            out.append((self.lam.arena, self.lam.garena))
        return out

    # ------------------------------------------------------------------------------------------------ optimizer
    def _segments(self):
        """(param slice, grad slice, lr index, weight decay, bf16 mirror or None) per optimizer group segment."""
        sp = self.splan
        segs = []
        bf = self.dt == torch.bfloat16
        segs.append((sp.arena[:sp.lin_total], sp.garena[:sp.lin_total], 0, self.wd, sp.lin16 if bf else None))
        if sp.la_off > sp.lin_total:
            segs.append((sp.arena[sp.lin_total:sp.la_off], sp.garena[sp.lin_total:sp.la_off], 0, self.wd, None))
        if sp.la_n:
            segs.append((sp.arena[sp.la_off:], sp.garena[sp.la_off:], 1, 0.0, None))
        segs.append((self.heads.arena, self.heads.garena, 0, self.wd, self.heads.a16 if bf else None))
        if self.lam is not None:
            segs.append((self.lam.arena[:2], self.lam.garena[:2], 2, 0.0, None))
        return segs

    def _optimizer_step(self):
        """clip_grad_norm_(max_norm) + AdamW over the arenas (lightning.py:200-238, distill.py:48)."""
        segs = self._segments()
        if self._opt_state is None:
            self._opt_state = [(ops.zeros(p.shape, torch.float32, self.device), ops.zeros(p.shape, torch.float32, self.device))
                               for p, _, _, _, _ in segs]
        ops.counter_inc(self.step_dev)
        use_clip = self.clip_norm > 0
        if use_clip:
            ops.memset(self.gnorm)
            for _, g, _, _, _ in segs:
                ops.sumsq(g, self.gnorm)
        for (p, g, li, wd, p16), (m, v) in zip(segs, self._opt_state):
            ops.adamw(p, g, m, v, self.lr_dev[li:li + 1], self.betas[0], self.betas[1], self.eps, wd, self.step_dev,
                      self.gnorm if use_clip else None, self.clip_norm, p16, True)
        # derived GEMM operands that are not plain casts (tap-major conv weights, weight-normalised pos conv)
        self.splan.rebuild_derived(self.dt, cast_lin=False)

    # ------------------------------------------------------------------------------------------------ data parallel
    # DDP's gradient mean (distill.py:41 strategy="ddp"): the Linear-weight gradients of transformer layer l (one
    # contiguous 28 MB slice of the gradient arena for HuBERT Base) are all-reduced as soon as the backward pass has
    # left that layer, on NCCL's own stream, while the earlier layers' backward keeps the SMs busy; everything else
    # (conv stack, biases, norms, log_alpha, heads, lambdas) follows after the backward. Works inside CUDA graph capture.
    def _reduce_async(self, t):
        import torch.distributed as dist

        if t.numel() == 0:
            return
        if t.is_cuda:
            w = dist.all_reduce(t, op=dist.ReduceOp.AVG, group=self.pg, async_op=True)
            self._pending.append((w, None))
        else:  # gloo (CPU tests): no AVG
            w = dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
            self._pending.append((w, t))

    def _layer_grads_ready(self, li):
        rng = self.splan.layer_lin_range[li]
        if rng is None:
            return
        lo, hi = rng
        side = self.splan.side_stream
        if side is not None:
            # the slice holds Linear-weight gradients only, all produced on the side stream: NCCL orders itself after it
            with torch.cuda.stream(side):
                self._reduce_async(self.splan.garena[lo:hi])
        else:
            self._reduce_async(self.splan.garena[lo:hi])
        self._covered.append((lo, hi))

    def _allreduce(self):
        if self.world <= 1:
            return
        g = self.splan.garena
        pos = 0
        for lo, hi in sorted(self._covered):
            self._reduce_async(g[pos:lo])
            pos = max(pos, hi)
        self._reduce_async(g[pos:])
        for _, ga in self.arenas()[1:]:
            self._reduce_async(ga)
        for w, t in self._pending:
            w.wait()
            if t is not None:
                t.mul_(1.0 / self.world)
        self._pending, self._covered = [], []

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
