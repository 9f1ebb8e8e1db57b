"""Tensor-core execution of ragged pruned students (the second training stage, final_distill.py:61-65, on the
architectures prune.py:33-59 writes: arbitrary conv channel counts and feed-forward widths, missing layers).

The tcgen05 / TMA path wants channel counts in multiples of 64.  Instead of specialising every kernel for ragged
widths, a pruned model gets a *padded shadow*: a structural copy whose conv channels and feed-forward widths are
rounded up to a multiple of 64 and whose extra rows / columns are zero.  Zero padding is closed under the whole
training step:
  * a padded conv channel has zero weight (and bias), so it outputs GELU(0) = 0; the next layer's weights for that
    input channel are zero, so nothing downstream sees it;
  * LayerNorms over a padded width normalise the real columns only (`valid_features` -> the kernels' d_valid);
  * gradients of padded entries are exact zeros (their upstream activations / downstream weights are zero), so
    AdamW leaves them at zero and the gradient norm is unchanged.
The real model's parameters (and their .grad) are narrow *views* of the shadow's, so `state_dict()`,
`load_state_dict()`, checkpoints and `get_num_params()` keep seeing the reference's shapes while every kernel runs
on the padded arena.  Attention needs nothing: head counts change, the head dimension (64) does not.
"""
import copy

import torch
from torch import nn

PAD = 64


def _up(n):
    return (n + PAD - 1) // PAD * PAD


def _has_gates(model):
    return any(n.endswith("log_alpha") for n, _ in model.named_parameters())


def is_ragged(model):
    """True if some conv channel count or feed-forward width is not a multiple of 64."""
    for blk in model.feature_extractor.conv_layers:
        if blk.conv.out_channels % PAD:
            return True
    for layer in model.encoder.transformer.layers:
        ff = layer.feed_forward
        if ff is not None and ff.intermediate_dense.out_features % PAD:
            return True
    return False


def _zeros_like_param(shape, ref):
    return nn.Parameter(torch.zeros(shape, dtype=ref.dtype, device=ref.device), requires_grad=ref.requires_grad)


def _pad_norm(ln, c, cp):
    """GroupNorm(c, c) / LayerNorm(c) -> the same over cp channels, in place (zero affine on the padding)."""
    if ln is None:
        return None
    ln.weight = _zeros_like_param((cp,), ln.weight)
    ln.bias = _zeros_like_param((cp,), ln.bias)
    if isinstance(ln, nn.GroupNorm):
        ln.num_groups = cp
        ln.num_channels = cp
    else:
        ln.normalized_shape = (cp,)
        ln.valid_features = c
    return ln


def _pad_linear(lin, out_p, in_p):
    new = nn.Linear(in_p, out_p, bias=lin.bias is not None, device=lin.weight.device, dtype=lin.weight.dtype)
    new.weight = _zeros_like_param((out_p, in_p), lin.weight)
    if lin.bias is not None:
        new.bias = _zeros_like_param((out_p,), lin.bias)
    return new


class Shadow:
    """Padded structural copy of `true` plus the aliasing of the true parameters onto it."""

    def __init__(self, true):
        if _has_gates(true):
            from .engine import DphShapeError

            raise DphShapeError(
                "bf16 tensor-core mode with HardConcrete gates needs conv channels / feed-forward widths in multiples "
                "of 64; use compute dtype float32 to keep pruning an already ragged model")
        self.true = true
        # the weight-normalised positional conv caches a derived (non-leaf) `weight`, which deepcopy refuses
        pconv = true.encoder.transformer.pos_conv_embed.conv
        cached = pconv.__dict__.pop("weight", None)
        try:
            self.model = m = copy.deepcopy(true)
        finally:
            if cached is not None:
                pconv.__dict__["weight"] = cached
        fe = m.feature_extractor
        cin_p = 1
        for blk in fe.conv_layers:
            conv = blk.conv
            c, cp = conv.out_channels, _up(conv.out_channels)
            new = nn.Conv1d(cin_p, cp, conv.kernel_size[0], stride=conv.stride[0], bias=conv.bias is not None,
                            device=conv.weight.device, dtype=conv.weight.dtype)
            new.weight = _zeros_like_param((cp, cin_p, conv.kernel_size[0]), conv.weight)
            if conv.bias is not None:
                new.bias = _zeros_like_param((cp,), conv.bias)
            blk.conv = new
            blk.layer_norm = _pad_norm(blk.layer_norm, c, cp)
            cin_p = cp
        fe.dummy_weight = _zeros_like_param((cin_p,), fe.dummy_weight)
        fp = m.encoder.feature_projection
        c = fp.projection.in_features
        fp.layer_norm = _pad_norm(fp.layer_norm, c, cin_p)
        fp.projection = _pad_linear(fp.projection, fp.projection.out_features, cin_p)
        for layer in m.encoder.transformer.layers:
            ff = layer.feed_forward
            if ff is None:
                continue
            f, d = ff.intermediate_dense.out_features, ff.intermediate_dense.in_features
            ff.intermediate_dense = _pad_linear(ff.intermediate_dense, _up(f), d)
            ff.output_dense = _pad_linear(ff.output_dense, d, _up(f))
        # copy the real values into the leading corner of every shadow parameter
        self.pairs = []
        shadow_params = dict(m.named_parameters())
        with torch.no_grad():
            for name, p in true.named_parameters():
                q = shadow_params[name]
                if tuple(q.shape) == tuple(p.shape):
                    q.copy_(p)
                else:
                    q[self._corner(p)].copy_(p)
                self.pairs.append((p, q))
        m.train(true.training)
        m.compute_dtype = true.compute_dtype

    @staticmethod
    def _corner(p):
        return tuple(slice(0, s) for s in p.shape)

    def alias(self):
        """Point the real parameters (and gradients) at the shadow's storage. Call after the shadow's Plan exists."""
        for p, q in self.pairs:
            p.data = q.data[self._corner(p)]
            if q.grad is not None:
                p.grad = q.grad[self._corner(p)]

    def aliased(self):
        for p, q in self.pairs:
            lo = q.data_ptr()
            if not (lo <= p.data_ptr() < lo + max(q.numel(), 1) * q.element_size()):
                return False
        return True

    def sync(self):
        """Mode / dtype follow the real model; gradients that user code dropped (zero_grad(set_to_none=True)) are
        re-attached."""
        m, t = self.model, self.true
        if m.training != t.training:
            m.train(t.training)
        m.compute_dtype = t.compute_dtype
        for p, q in self.pairs:
            if q.requires_grad != p.requires_grad:
                q.requires_grad_(p.requires_grad)
            if q.grad is not None and (p.grad is None or p.grad.data_ptr() != q.grad.data_ptr()):
                if p.grad is None:
                    q.grad.zero_()
                p.grad = q.grad[self._corner(p)]
