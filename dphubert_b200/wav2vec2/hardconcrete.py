"""HardConcrete (L0) gate — host-side mirror of the reference module `wav2vec2/hardconcrete.py:14-122`.

Same constructor, attributes (`log_alpha`, `beta`, `limit_l`, `limit_r`, `bias`, `eps`, `compiled_mask`) and
semantics.  Train-mode sampling and `l0_norm()` run as CUDA kernels (`dph_hardconcrete_*`) when the
parameter lives on a GPU; the eval-mode "compiled" mask is the cold, bit-exact top-k path used by `prune()`
and stays on whatever device the parameter is on (the reference prunes on CPU, `prune.py:11-32`).

Extension over the reference: `noise` — if set to a tensor of uniforms in (eps, 1-eps), the next train-mode
forward uses it instead of drawing (parity tests feed both sides the same `u`).
"""
import math

import torch
import torch.nn as nn


def _no_cpu(what):
    raise RuntimeError(
        "dphubert_b200: %s needs the CUDA kernels and a GPU-resident module (there is no CPU fallback)" % what
    )


class HardConcrete(nn.Module):
    def __init__(self, n_in, init_mean=0.5, init_std=0.01, temperature=2 / 3, stretch=0.1, eps=1e-6):
        super().__init__()
        self.n_in = n_in
        self.limit_l = -stretch
        self.limit_r = 1.0 + stretch
        self.log_alpha = nn.Parameter(torch.zeros(n_in))
        self.beta = temperature
        self.init_mean = init_mean
        self.init_std = init_std
        # shift that turns sigmoid(log_alpha + bias) into P(gate != 0)  (hardconcrete.py:64)
        self.bias = -self.beta * math.log(-self.limit_l / self.limit_r)
        self.eps = eps
        self.compiled_mask = None
        self.noise = None
        self.reset_parameters()

    def reset_parameters(self):
        self.compiled_mask = None
        mean = math.log(1 - self.init_mean) - math.log(self.init_mean)
        self.log_alpha.data.normal_(mean, self.init_std)

    # ------------------------------------------------------------------ expected L0
    def l0_norm(self):
        """Expected number of open gates, sum(sigmoid(log_alpha + bias))  (hardconcrete.py:76-83)."""
        if self.log_alpha.is_cuda:
            from .. import ops

            return ops.l0_norm(self.log_alpha, self.bias)
        if self.training:
            _no_cpu("HardConcrete.l0_norm() in training mode")
        # eval-mode host path: cold, only feeds prune()'s mask compilation and offline size reports
        return (self.log_alpha.detach() + self.bias).sigmoid().sum()

    # ------------------------------------------------------------------ mask
    def _draw_noise(self):
        if self.noise is not None:
            u = self.noise.to(self.log_alpha.device, torch.float32)
            self.noise = None
            return u
        return self.log_alpha.new_empty(self.n_in).uniform_(self.eps, 1 - self.eps)

    def forward(self):
        if self.training:
            self.compiled_mask = None
            u = self._draw_noise()
            if self.log_alpha.is_cuda:
                from .. import ops

                return ops.hardconcrete_sample(self.log_alpha, u, self.beta, self.limit_l, self.limit_r)
            _no_cpu("HardConcrete sampling")
        if self.compiled_mask is None:
            self.compiled_mask = self._compile_mask()
        return self.compiled_mask

    def _compile_mask(self):
        """Deterministic mask: zero the round(expected #zeros) smallest gates, keep the soft value of the rest
        (hardconcrete.py:101-114; python round() = banker's rounding; the 0.8 factor is the reference's)."""
        la = self.log_alpha.detach()
        expected_open = (la + self.bias).sigmoid().sum().item()
        num_zeros = round(self.n_in - expected_open)
        soft = torch.sigmoid(la / self.beta * 0.8)
        if num_zeros > 0:
            _, idx = torch.topk(soft, k=num_zeros, largest=False)
            soft[idx] = 0.0
        return soft

    def extra_repr(self):
        return str(self.n_in)

    def __repr__(self):
        return "{}({})".format(self.__class__.__name__, self.extra_repr())
