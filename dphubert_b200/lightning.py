"""Training-side drop-in surface — mirror of the reference `lightning.py` (DistillModule, DistillLoss, the two LR
schedulers) without a pytorch_lightning dependency.

`DistillModule` keeps the reference's constructor keywords (lightning.py:143-164), attribute names, state-dict prefixes
(`teacher_model.`, `student_model.`, `distill_linear_projs.N.`, `lambda1`, `lambda2`) and `_step` / `training_step` /
`validation_step` / `configure_optimizers` entry points.  `training_step` returns the differentiable loss exactly like
the reference (lightning.py:298-300) for an external optimizer loop; `native_training_step` is the fused native step
(`dphubert_b200.trainer.NativeDistillStep`: forward, loss, backward, all-reduce, clip, AdamW and the LR schedule in one
launch sequence / CUDA graph) and refuses to be mixed with the former.  `train_dataloader` / `val_dataloader`
(lightning.py:306-342) iterate the HBM-resident waveform pool of `dphubert_b200.data`.
"""
import math
import pathlib
from typing import List, Optional, Union

import torch
import torch.nn as nn
from torch.optim.optimizer import Optimizer

from .wav2vec2.model import Wav2Vec2Model


class LinearDecayLRScheduler(torch.optim.lr_scheduler.LRScheduler):
    """Linear warm-up then linear decay to zero, keyed on `_step_count` (reference lightning.py:22-44; the `verbose`
    kwarg the reference forwards no longer exists in torch >= 2.x and is dropped)."""

    def __init__(self, optimizer: Optimizer, warmup_updates: int, max_updates: int, last_epoch: int = -1):
        self.warmup_updates = warmup_updates
        self.max_updates = max_updates
        super().__init__(optimizer, last_epoch=last_epoch)

    def get_lr(self):
        if self._step_count <= self.warmup_updates:
            return [self._step_count / self.warmup_updates * base_lr for base_lr in self.base_lrs]
        elif self._step_count >= self.max_updates:
            return [0.0 for _ in self.base_lrs]
        pct_remaining = (self.max_updates - self._step_count) / (self.max_updates - self.warmup_updates)
        return [base_lr * pct_remaining for base_lr in self.base_lrs]


class TriStageLRScheduler(torch.optim.lr_scheduler.LRScheduler):
    """Warm-up, hold, exponential decay (reference lightning.py:47-88)."""

    def __init__(self, optimizer: Optimizer, warmup_updates: int, hold_updates: int, decay_updates: int,
                 init_lr_scale: float = 0.01, final_lr_scale: float = 0.05, last_epoch: int = -1):
        self.warmup_updates = warmup_updates
        self.hold_updates = hold_updates
        self.decay_updates = decay_updates
        self.init_lr_scale = init_lr_scale
        self.final_lr_scale = final_lr_scale
        super().__init__(optimizer, last_epoch=last_epoch)

    def get_lr(self):
        n = self._step_count
        if n <= self.warmup_updates:
            return [b * (self.init_lr_scale + n / self.warmup_updates * (1 - self.init_lr_scale)) for b in self.base_lrs]
        if n <= self.warmup_updates + self.hold_updates:
            return list(self.base_lrs)
        if n <= self.warmup_updates + self.hold_updates + self.decay_updates:
            f = math.exp(math.log(self.final_lr_scale) * (n - self.warmup_updates - self.hold_updates) / self.decay_updates)
            return [b * f for b in self.base_lrs]
        return [b * self.final_lr_scale for b in self.base_lrs]


class DistillLoss(nn.Module):
    """w2 * MSE + w1 * L1 + wc * (-cos) over (batch, layer, time, feature) tensors (reference lightning.py:91-139).
    On CUDA tensors the three reductions and their backward are the fused `dph_distill_loss_*` kernels."""

    def __init__(self, l2_weight, l1_weight, cos_weight, cos_type):
        super().__init__()
        self.l2_weight = l2_weight
        self.l1_weight = l1_weight
        self.cos_weight = cos_weight
        self.cos_type = cos_type
        assert cos_type in ["raw", "log_sig"], cos_type

    def __repr__(self) -> str:
        return "{}(l2={}, l1={}, {}_cos={})".format(self.__class__.__name__, self.l2_weight, self.l1_weight,
                                                    self.cos_type, self.cos_weight)

    def forward(self, input: torch.Tensor, target: torch.Tensor):
        from . import _lib, engine

        if not engine._is_native(input):
            raise _lib.DphError("DistillLoss runs only on CUDA tensors through the native kernels (no CPU fallback)")
        total, parts = _DistillLossFn.apply(input, target, float(self.l2_weight), float(self.l1_weight),
                                            float(self.cos_weight), 0 if self.cos_type == "raw" else 1)
        mse = parts[1] if self.l2_weight != 0 else 0
        l1 = parts[2] if self.l1_weight != 0 else 0
        cos = parts[3] if self.cos_weight != 0 else 0
        return total, (mse, l1, cos)


class _DistillLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, w2, w1, wc, cos_type):
        from . import ops

        a_c = a.detach().contiguous()
        b_c = b.detach().to(a_c.dtype).contiguous()
        D = a_c.shape[-1]
        rows = a_c.numel() // D
        dev = a_c.device
        rowstat = torch.empty(rows * 3, dtype=torch.float32, device=dev)
        acc = torch.empty(4, dtype=torch.float32, device=dev)
        out = torch.empty(4, dtype=torch.float32, device=dev)
        ops.distill_loss_fwd(a_c, b_c, rowstat, acc, out, rows, D, w2, w1, wc, cos_type, True, True, rows)
        ctx.save_for_backward(a_c, b_c, rowstat)
        ctx.hp = (w2, w1, wc, cos_type, rows, D)
        ctx.mark_non_differentiable(out)
        return out[0].clone(), out

    @staticmethod
    def backward(ctx, g, _g_parts):
        from . import ops

        a_c, b_c, rowstat = ctx.saved_tensors
        w2, w1, wc, cos_type, rows, D = ctx.hp
        da = torch.empty_like(a_c)
        ops.distill_loss_bwd(a_c, b_c, rowstat, g.reshape(1).float().contiguous(), da, rows, rows, D, w2, w1, wc, cos_type)
        return da, None, None, None, None, None


class DistillModule(nn.Module):
    """Joint distillation + structured pruning module (reference lightning.py:142-304)."""

    def __init__(self, *, teacher_model: Wav2Vec2Model, student_model: Wav2Vec2Model, distill_mode: str,
                 distill_layers: List[int], distill_linear_projs: nn.ModuleList, distill_loss: DistillLoss,
                 learning_rate: float, weight_decay: float, warmup_updates: int, max_updates: int, use_reg: bool,
                 reg_learning_rate: Optional[float], target_sparsity: Optional[float],
                 sparsity_warmup_updates: Optional[int], tsv_dir: Union[str, pathlib.Path] = ".",
                 train_subset: str = "", seconds_per_batch: float = 0.0, num_workers: int = 0,
                 gradient_clip_val: float = 10.0):
        super().__init__()
        self.teacher_model = teacher_model
        self.student_model = student_model
        self.original_num_params = sum(p.numel() for p in teacher_model.parameters())
        assert distill_mode in ["layer2layer", "predlayer"], distill_mode
        assert len(distill_layers) == len(distill_linear_projs)
        self.distill_mode = distill_mode
        self.distill_layers = distill_layers
        self.distill_linear_projs = distill_linear_projs
        self.distill_loss = distill_loss
        self.learning_rate = learning_rate
        self.weight_decay = weight_decay
        self.warmup_updates = warmup_updates
        self.max_updates = max_updates
        self.use_reg = use_reg
        self.reg_learning_rate = reg_learning_rate
        self.target_sparsity = target_sparsity
        self.sparsity_warmup_updates = sparsity_warmup_updates
        if self.use_reg:
            self.lambda1 = nn.Parameter(torch.tensor(0.0))
            self.lambda2 = nn.Parameter(torch.tensor(0.0))
        self.tsv_dir, self.train_subset = tsv_dir, train_subset
        self.seconds_per_batch, self.num_workers = seconds_per_batch, num_workers
        self.gradient_clip_val = gradient_clip_val  # distill.py:48 (a Trainer argument in the reference)
        self.global_step = 0
        self.current_epoch = 0
        self.logged = {}
        self._native = None

    # ---- Lightning-compatible conveniences
    def log_dict(self, d, **kwargs):
        self.logged.update(d)

    def configure_optimizers(self):
        """Same three parameter groups as the reference (lightning.py:200-238) for callers that drive `_step` with a
        torch optimizer; `training_step` uses the fused native AdamW instead."""
        main_params = [p for n, p in self.student_model.named_parameters() if "log_alpha" not in n and p.requires_grad]
        main_params.extend(list(self.distill_linear_projs.parameters()))
        pgs = [{"params": main_params, "lr": self.learning_rate, "weight_decay": self.weight_decay, "name": "main_params"}]
        if self.use_reg:
            pgs.extend([
                {"params": [p for n, p in self.student_model.named_parameters() if "log_alpha" in n],
                 "lr": self.reg_learning_rate, "weight_decay": 0.0, "name": "log_alpha"},
                {"params": [self.lambda1, self.lambda2], "lr": -self.reg_learning_rate, "weight_decay": 0.0,
                 "name": "lambda"},
            ])
        # torch >= 2 validates lr >= 0 in the constructor; the reference's negative-lr group is set afterwards
        for pg in pgs:
            pg["_lr"] = pg["lr"]
            pg["lr"] = abs(pg["lr"])
        optimizer = torch.optim.AdamW(pgs)
        for pg in optimizer.param_groups:
            pg["lr"] = pg.pop("_lr")
            pg["initial_lr"] = pg["lr"]
        lr_scheduler = LinearDecayLRScheduler(optimizer, warmup_updates=self.warmup_updates, max_updates=self.max_updates)
        return {"optimizer": optimizer, "lr_scheduler": {"scheduler": lr_scheduler, "interval": "step"}}

    def _get_target_sparsity(self):
        if self.global_step >= self.sparsity_warmup_updates:
            return self.target_sparsity
        return self.target_sparsity * (self.global_step / self.sparsity_warmup_updates)

    def _step(self, batch, batch_idx, mode):
        """Loss under autograd (reference lightning.py:245-296): native forward/backward of both models through the
        autograd bridge, heads as torch Linears, fused loss kernel."""
        waveforms, lengths = batch
        self.teacher_model.eval()
        with torch.no_grad():
            teacher_hiddens, _ = self.teacher_model.extract_features(waveforms, lengths)
            teacher_hiddens = torch.stack([teacher_hiddens[idx] for idx in self.distill_layers], dim=1)
        student_hiddens, _ = self.student_model.extract_features(waveforms, lengths)
        new_student_hiddens = []
        for idx, proj in zip(self.distill_layers, self.distill_linear_projs):
            if self.distill_mode == "layer2layer":
                src = student_hiddens[idx]
            elif self.distill_mode == "predlayer":
                src = student_hiddens[-1]
            else:
                raise ValueError(f"Invalid distill mode: {self.distill_mode}")
            new_student_hiddens.append(proj(src.float()))
        student_hiddens = torch.stack(new_student_hiddens, dim=1)
        loss_distill, (loss_mse, loss_l1, loss_cos) = self.distill_loss(student_hiddens, teacher_hiddens.float())
        if self.use_reg:
            cur_target_sparsity = self._get_target_sparsity()
            cur_expected_sparsity = 1.0 - self.student_model.get_num_params() / self.original_num_params
            loss_reg = self.lambda1 * (cur_expected_sparsity - cur_target_sparsity) \
                + self.lambda2 * (cur_expected_sparsity - cur_target_sparsity) ** 2
        else:
            loss_reg = 0
        loss = loss_distill + loss_reg
        self.log_dict({f"{mode}_loss": loss, f"{mode}_loss_distill": loss_distill, f"{mode}_loss_mse": loss_mse,
                       f"{mode}_loss_l1": loss_l1, f"{mode}_loss_cos": loss_cos, f"{mode}_loss_reg": loss_reg})
        if mode == "train" and self.use_reg:
            self.log_dict({"sparsity_expected": cur_expected_sparsity, "sparsity_target": cur_target_sparsity,
                           "lambda1": self.lambda1, "lambda2": self.lambda2})
        return loss

    def native_step(self):
        """The fused training step bound to this module's models / heads / lambdas / hyper-parameters."""
        if self._native is None:
            from .trainer import NativeDistillStep

            dl = self.distill_loss
            self._native = NativeDistillStep(
                self.teacher_model, self.student_model, self.distill_layers, list(self.distill_linear_projs),
                self.distill_mode, dl.l2_weight, dl.l1_weight, dl.cos_weight, dl.cos_type, self.use_reg,
                [self.lambda1, self.lambda2] if self.use_reg else None, self.original_num_params, self.learning_rate,
                self.weight_decay, self.reg_learning_rate, self.warmup_updates, self.max_updates, self.target_sparsity,
                self.sparsity_warmup_updates, self.gradient_clip_val)
            self._native.global_step = self.global_step
        return self._native

    def training_step(self, batch, batch_idx):
        """Reference contract (lightning.py:298-300): the differentiable loss of `_step`, for an external optimizer
        loop (`pl.Trainer`, or `configure_optimizers()` driven by hand).  Refuses to run once the fused native step
        owns the optimisation of this module — mixing the two would apply two updates per batch."""
        if self._native is not None and self._native.global_step > 0:
            raise RuntimeError(
                "DistillModule.training_step: this module is being optimised by the fused native step "
                "(native_training_step / native_step()); an external optimizer loop on top would double-step. "
                "Use one or the other.")
        return self._step(batch, batch_idx, mode="train")

    def native_training_step(self, batch, batch_idx=0):
        """One full optimisation step (forward, backward, all-reduce, clip, AdamW, schedule) as a single launch
        sequence — returns the loss terms on the device as in `NativeDistillStep.run_step`.  This replaces
        `training_step` + Lightning's optimizer loop; it must not be called from inside one."""
        waveforms, lengths = batch
        step = self.native_step()
        out = step.step(waveforms, lengths)   # eager the first time a batch shape is seen, a replayed graph afterwards
        self.global_step = step.global_step
        return out

    # ---- Lightning checkpoint hooks (trainer.fit(ckpt_path=...), distill.py:141-144): the native optimizer state
    def on_save_checkpoint(self, checkpoint):
        if self._native is not None:
            checkpoint["native_step"] = self._native.state_dict()
        checkpoint["global_step"] = self.global_step

    def on_load_checkpoint(self, checkpoint):
        self.global_step = int(checkpoint.get("global_step", self.global_step))
        if "native_step" in checkpoint:
            self.native_step().load_state_dict(checkpoint["native_step"])
            self.global_step = self._native.global_step

    # ---- data (lightning.py:306-342): same samplers / collate policy, batches assembled on the device
    def _pool_and_lengths(self, subset, min_len=32000, max_len=250000):
        from . import data

        cache = self.__dict__.setdefault("_pools", {})
        if subset not in cache:
            paths, _, lens = data.read_manifest(self.tsv_dir, subset)
            keep = [i for i, n in enumerate(lens) if min_len <= n <= max_len]   # what the bucket sampler can emit
            device = next(self.student_model.parameters()).device
            cache[subset] = (data.load_pool(paths, lens, device, keep=keep, workers=max(1, int(self.num_workers or 1))),
                             lens)
        return cache[subset]

    def _bucket_sampler(self, lens):
        from . import data

        return data.BucketizeBatchSampler(lens, num_buckets=1000, max_token_count=self.seconds_per_batch * 16000,
                                          min_len=32000, max_len=250000, shuffle=False)

    def train_dataloader(self, quantum: int = 0):
        """lightning.py:306-324: 1000 length buckets, <= seconds_per_batch of audio per batch, 2 s <= len <= 15.6 s,
        whole batches dealt to ranks with a seeded shuffle, crop to the shortest utterance with random offsets."""
        import torch.distributed as dist

        from . import data

        pool, lens = self._pool_and_lengths(self.train_subset)
        sampler = self._bucket_sampler(lens)
        if dist.is_available() and dist.is_initialized():
            sampler = data.DistributedBatchSampler(sampler, shuffle=True)
        else:
            sampler = data.DistributedBatchSampler(sampler, num_replicas=1, rank=0, shuffle=True)
        sampler.set_epoch(self.current_epoch)
        return data.DeviceLoader(pool, sampler, pad=False, rand_crop=True, quantum=quantum)

    def val_dataloader(self):
        """lightning.py:326-342 (not distributed: every rank sees the whole validation set)."""
        from . import data

        pool, lens = self._pool_and_lengths("valid")
        return data.DeviceLoader(pool, self._bucket_sampler(lens), pad=False, rand_crop=True)

    def validation_step(self, batch, batch_idx):
        with torch.no_grad():
            return self._step(batch, batch_idx, mode="valid")
