"""CPU oracle for the DPHuBERT joint distill-and-prune step — TEST INFRASTRUCTURE ONLY.

A functional fp32 restatement (plain torch CPU ops over a reference-format state dict) of the reference algorithm
on the hot path.  It is imported only by `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline /
`--impl reference` legs; the product (`dphubert_b200/`) never imports it and has no CPU fallback.

Parity pinning: the reference (seas2nada/DPHuBERT) ships no tests and no golden vectors, so this file
is pinned against the reference *itself*: `oracle/make_golden.py` imports `/root/reference/wav2vec2` in the build
container, runs both on the same seeded inputs and commits the reference's outputs under `tests/golden/`;
`tests/test_oracle.py` replays them without the reference being present.

Each function cites the reference file:line it restates (paths relative to the reference checkout).
"""
import math
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# --------------------------------------------------------------------------------------------------------------
# HardConcrete gates (wav2vec2/hardconcrete.py)
# --------------------------------------------------------------------------------------------------------------
HC_BETA = 2.0 / 3.0       # hardconcrete.py:31 temperature
HC_L, HC_R = -0.1, 1.1    # hardconcrete.py:58-59 stretch limits
HC_EPS = 1e-6
HC_BIAS = -HC_BETA * math.log(-HC_L / HC_R)  # hardconcrete.py:64


def hc_sample(log_alpha: Tensor, u: Tensor) -> Tensor:
    """Training-mode mask for given uniform noise u in (eps, 1-eps)  — hardconcrete.py:96-99."""
    s = torch.sigmoid((torch.log(u / (1.0 - u)) + log_alpha) / HC_BETA)
    return torch.clamp(s * (HC_R - HC_L) + HC_L, 0.0, 1.0)


def hc_l0(log_alpha: Tensor) -> Tensor:
    """Expected number of non-zero gates — hardconcrete.py:76-83."""
    return torch.sigmoid(log_alpha + HC_BIAS).sum()


def hc_eval_mask(log_alpha: Tensor) -> Tensor:
    """Deterministic eval-mode mask — hardconcrete.py:101-114 (python round() = banker's rounding)."""
    n = log_alpha.numel()
    num_zeros = round(n - hc_l0(log_alpha).item())
    soft = torch.sigmoid(log_alpha.detach() / HC_BETA * 0.8)
    if num_zeros > 0:
        _, idx = torch.topk(soft, k=num_zeros, largest=False)
        soft = soft.clone()
        soft[idx] = 0.0
    return soft


# --------------------------------------------------------------------------------------------------------------
# model forward (wav2vec2/components.py, wav2vec2/model.py)
# --------------------------------------------------------------------------------------------------------------
def _conv_cfg(cfg):
    c = cfg.get("extractor_conv_layer_config")
    return c if c is not None else [(512, 10, 5)] + [(512, 3, 2)] * 4 + [(512, 2, 2)] * 2


def is_wavlm(cfg) -> bool:
    return "encoder_remaining_heads" in cfg  # model.py:175


def feature_extractor(sd, cfg, wave: Tensor, lengths: Optional[Tensor], masks: Dict[str, Tensor]):
    """components.py:158-185 with ConvLayerBlock.forward components.py:94-120."""
    x = wave.unsqueeze(1)
    mode = cfg["extractor_mode"]
    for i, (cout, k, s) in enumerate(_conv_cfg(cfg)):
        pre = "feature_extractor.conv_layers.%d." % i
        x = F.conv1d(x, sd[pre + "conv.weight"], sd.get(pre + "conv.bias"), stride=s)
        if mode == "group_norm" and i == 0:
            x = F.group_norm(x, cout, sd[pre + "layer_norm.weight"], sd[pre + "layer_norm.bias"])
        elif mode == "layer_norm":
            x = F.layer_norm(x.transpose(-2, -1), (cout,), sd[pre + "layer_norm.weight"],
                             sd[pre + "layer_norm.bias"]).transpose(-2, -1)  # components.py:54-61
        x = F.gelu(x)
        m = masks.get(pre + "hard_concrete")
        if m is not None:
            x = x * m.unsqueeze(-1)  # components.py:112-114
        if lengths is not None:
            lengths = torch.div(lengths - k, s, rounding_mode="floor") + 1
            lengths = torch.max(torch.zeros_like(lengths), lengths)  # components.py:116-119
    x = x.transpose(1, 2) * sd["feature_extractor.dummy_weight"]  # components.py:183-184
    return x, lengths


def _rel_bucket(rel: Tensor, num_buckets: int, max_distance: int) -> Tensor:
    """components.py:565-600 (bidirectional)."""
    nb = num_buckets // 2
    buckets = (rel > 0).to(torch.long) * nb
    rp = rel.abs()
    max_exact = nb // 2
    is_small = rp < max_exact
    large = max_exact + (torch.log(rp.float() / max_exact) / math.log(max_distance / max_exact)
                         * (nb - max_exact)).to(torch.long)
    large = torch.min(large, torch.full_like(large, nb - 1))
    return buckets + torch.where(is_small, rp, large)


def _attention(sd, cfg, pre, x, add_mask, H, Dh, masks, pos_bias, layer_idx):
    """SelfAttention.forward components.py:379-436; WavLMSelfAttention.forward components.py:604-659."""
    B, T, D = x.shape
    new_pos_bias = pos_bias
    if is_wavlm(cfg):
        Htot = cfg["encoder_total_num_heads"][layer_idx]
        remaining = cfg["encoder_remaining_heads"][layer_idx]
        if pos_bias is None:
            emb = sd["encoder.transformer.layers.0.attention.rel_attn_embed.weight"]  # components.py:546-563
            ctx = torch.arange(T)[:, None]
            mem = torch.arange(T)[None, :]
            bucket = _rel_bucket(mem - ctx, cfg["encoder_num_buckets"], cfg["encoder_max_distance"])
            pos_bias = emb[bucket].permute(2, 0, 1).unsqueeze(0).repeat(B, 1, 1, 1).reshape(B * Htot, T, T)
            new_pos_bias = pos_bias
        ql = x.view(B, T, Htot, -1).permute(0, 2, 1, 3)  # components.py:637-644
        g = F.linear(ql, sd[pre + "gru_rel_pos_linear.weight"], sd[pre + "gru_rel_pos_linear.bias"])
        ga, gb = torch.sigmoid(g.view(B, Htot, T, 2, 4).sum(-1)).chunk(2, dim=-1)
        gate = ga * (gb * sd[pre + "gru_rel_pos_const"] - 1.0) + 2.0
        bias = gate.view(B * Htot, -1, 1) * pos_bias
        bias = bias.reshape(B, Htot, T, T)[:, remaining, :, :]
        add_mask = bias if add_mask is None else bias + add_mask
    q = F.linear(x, sd[pre + "q_proj.weight"], sd[pre + "q_proj.bias"]).view(B, T, H, Dh).transpose(2, 1)
    k = F.linear(x, sd[pre + "k_proj.weight"], sd[pre + "k_proj.bias"]).view(B, T, H, Dh).permute(0, 2, 3, 1)
    v = F.linear(x, sd[pre + "v_proj.weight"], sd[pre + "v_proj.bias"]).view(B, T, H, Dh).transpose(2, 1)
    w = (Dh ** -0.5 * q) @ k
    if add_mask is not None:
        w = w + add_mask
    w = w - w.max(dim=-1, keepdim=True)[0]
    w = torch.softmax(w, dim=-1)
    o = w @ v
    hm = masks.get(pre + "hard_concrete_for_heads")
    if hm is not None:
        o = o * hm.unsqueeze(-1).unsqueeze(-1)
    o = o.transpose(2, 1).reshape(B, T, H * Dh)
    o = F.linear(o, sd[pre + "out_proj.weight"], sd[pre + "out_proj.bias"])
    lm = masks.get(pre + "hard_concrete_for_layer")
    if lm is not None:
        o = o * lm
    return o, new_pos_bias


def _feed_forward(sd, pre, x, masks):
    """FeedForward.forward components.py:726-748 (dropout off)."""
    x = F.gelu(F.linear(x, sd[pre + "intermediate_dense.weight"], sd[pre + "intermediate_dense.bias"]))
    im = masks.get(pre + "hard_concrete_for_intermediate")
    if im is not None:
        x = x * im
    x = F.linear(x, sd[pre + "output_dense.weight"], sd[pre + "output_dense.bias"])
    lm = masks.get(pre + "hard_concrete_for_layer")
    if lm is not None:
        x = x * lm
    return x


def _ln(x, sd, pre):
    return F.layer_norm(x, (x.shape[-1],), sd[pre + "weight"], sd[pre + "bias"])


def _pos_conv(sd, cfg, x):
    """ConvolutionalPositionalEmbedding.forward components.py:319-333 (weight_norm dim=2: components.py:298-307)."""
    pre = "encoder.transformer.pos_conv_embed.conv."
    v, g = sd[pre + "weight_v"], sd[pre + "weight_g"]
    w = g * v / torch.linalg.vector_norm(v, dim=(0, 1), keepdim=True)
    K = cfg["encoder_pos_conv_kernel"]
    y = F.conv1d(x.transpose(-2, -1), w, sd[pre + "bias"], padding=K // 2, groups=cfg["encoder_pos_conv_groups"])
    if K % 2 == 0:
        y = y[..., :-1]
    return F.gelu(y).transpose(-2, -1)


def extract_features(sd, cfg, wave: Tensor, lengths: Optional[Tensor] = None, masks: Optional[Dict] = None,
                     num_layers: Optional[int] = None):
    """Wav2Vec2Model.extract_features model.py:57-107 → Encoder.extract_features components.py:995-1003 →
    Transformer.get_intermediate_outputs components.py:909-927. Dropout disabled (parity configuration).
    `masks` maps '<module path>.hard_concrete…' (no trailing '.log_alpha') to the mask tensor."""
    masks = masks or {}
    if cfg["normalize_waveform"]:  # model.py:96-103
        if lengths is not None:
            ws = [F.layer_norm(w[:n], (int(n),)) for w, n in zip(wave, lengths)]
            wave = torch.nn.utils.rnn.pad_sequence(ws, batch_first=True)
        else:
            wave = F.layer_norm(wave, wave.shape[-1:])
    x, lengths = feature_extractor(sd, cfg, wave, lengths, masks)
    # Encoder._preprocess components.py:968-984
    x = _ln(x, sd, "encoder.feature_projection.layer_norm.")
    x = F.linear(x, sd["encoder.feature_projection.projection.weight"], sd["encoder.feature_projection.projection.bias"])
    add_mask = None
    if lengths is not None:
        B, T, _ = x.shape
        pad = torch.arange(T).expand(B, T) >= lengths[:, None]
        x = x.masked_fill(pad.unsqueeze(-1), 0.0)
        add_mask = (-10000.0 * pad[:, None, None, :].to(x.dtype)).expand(B, 1, T, T)
    hiddens = [x]
    # Transformer._preprocess components.py:885-892 (layer_norm_first of the Transformer = NOT config flag, :1283)
    x = x + _pos_conv(sd, cfg, x)
    pre_norm = cfg["encoder_layer_norm_first"]
    if not pre_norm:
        x = _ln(x, sd, "encoder.transformer.layer_norm.")
    pos_bias = None
    Dh = cfg.get("encoder_head_dim")
    for l in range(cfg["encoder_num_layers"]):
        pre = "encoder.transformer.layers.%d." % l
        if cfg["encoder_use_attention"][l]:  # EncoderLayer.forward components.py:814-857
            if is_wavlm(cfg):
                H = len(cfg["encoder_remaining_heads"][l])
                dh = cfg["encoder_embed_dim"] // cfg["encoder_total_num_heads"][l]
            else:
                H, dh = cfg["encoder_num_heads"][l], Dh
            res = x
            xa = _ln(x, sd, pre + "layer_norm.") if pre_norm else x
            a, pos_bias = _attention(sd, cfg, pre + "attention.", xa, add_mask, H, dh, masks, pos_bias, l)
            x = res + a
        if pre_norm:
            if cfg["encoder_use_feed_forward"][l]:
                x = x + _feed_forward(sd, pre + "feed_forward.", _ln(x, sd, pre + "final_layer_norm."), masks)
        else:
            x = _ln(x, sd, pre + "layer_norm.")
            if cfg["encoder_use_feed_forward"][l]:
                x = x + _feed_forward(sd, pre + "feed_forward.", x, masks)
            x = _ln(x, sd, pre + "final_layer_norm.")
        hiddens.append(x)
        if num_layers is not None and len(hiddens) - 1 >= num_layers:
            break
    return hiddens, lengths


# --------------------------------------------------------------------------------------------------------------
# expected size (get_num_params tree) and the training loss (lightning.py)
# --------------------------------------------------------------------------------------------------------------
def expected_num_params(sd, cfg):
    """Wav2Vec2Model.get_num_params model.py:109-113 and the per-module formulas components.py:122-134, 187-196,
    276-277, 438-449, 750-761, 859-865, 929-934, 1005-1009. Differentiable wrt every log_alpha in `sd`."""
    def l0(key, static):
        la = sd.get(key + ".log_alpha")
        return hc_l0(la) if la is not None else static

    in_ch = 1
    total = 0
    for i, (cout, k, s) in enumerate(_conv_cfg(cfg)):
        pre = "feature_extractor.conv_layers.%d." % i
        out = l0(pre + "hard_concrete", cout)
        n = in_ch * out * k
        if (pre + "conv.bias") in sd:
            n = n + out
        if (pre + "layer_norm.weight") in sd:
            n = n + out * 2
        total = total + n
        in_ch = out
    total = total + in_ch  # dummy weight, components.py:194
    D = cfg["encoder_embed_dim"]
    total = total + in_ch * 2 + (in_ch + 1) * D  # components.py:276-277
    pc = "encoder.transformer.pos_conv_embed.conv."
    total = total + sd[pc + "weight_v"].numel() + sd[pc + "weight_g"].numel() + sd[pc + "bias"].numel() + D * 2
    for l in range(cfg["encoder_num_layers"]):
        pre = "encoder.transformer.layers.%d." % l
        total = total + D * 2 * 2
        if cfg["encoder_use_attention"][l]:
            if is_wavlm(cfg):
                H = len(cfg["encoder_remaining_heads"][l])
                dh = D // cfg["encoder_total_num_heads"][l]
            else:
                H, dh = cfg["encoder_num_heads"][l], cfg["encoder_head_dim"]
            h = l0(pre + "attention.hard_concrete_for_heads", H)
            n = (D + 1) * h * dh * 3 + (h * dh + 1) * D
            g = sd.get(pre + "attention.hard_concrete_for_layer.log_alpha")
            if g is not None:
                n = n * hc_l0(g)
            total = total + n
        if cfg["encoder_use_feed_forward"][l]:
            f = l0(pre + "feed_forward.hard_concrete_for_intermediate", cfg["encoder_ff_interm_features"][l])
            n = (D + 1) * f + (f + 1) * D
            g = sd.get(pre + "feed_forward.hard_concrete_for_layer.log_alpha")
            if g is not None:
                n = n * hc_l0(g)
            total = total + n
    return total


def distill_loss(pred: Tensor, target: Tensor, l2_weight, l1_weight, cos_weight, cos_type="raw"):
    """DistillLoss.forward lightning.py:116-139. pred/target: (batch, layer, time, feature)."""
    mse = F.mse_loss(pred, target) if l2_weight != 0 else 0
    l1 = F.l1_loss(pred, target) if l1_weight != 0 else 0
    cos = 0
    if cos_weight != 0:
        c = F.cosine_similarity(pred, target, dim=-1)
        cos = -c.mean() if cos_type == "raw" else -c.sigmoid().log().mean()
    return l2_weight * mse + l1_weight * l1 + cos_weight * cos, (mse, l1, cos)


def gate_keys(sd) -> List[str]:
    return [k[: -len(".log_alpha")] for k in sd if k.endswith(".log_alpha")]


def distill_step_loss(t_sd, t_cfg, s_sd, s_cfg, heads, distill_layers, wave, lengths, noise=None, lambdas=None,
                      target_sparsity=None, original_num_params=None, weights=(0.0, 1.0, 1.0), cos_type="raw",
                      mode="layer2layer"):
    """DistillModule._step lightning.py:245-296 (dropout off). `heads`: list of (weight, bias) per distill layer;
    `noise`: {gate key: uniform noise}; returns (loss, dict of terms)."""
    with torch.no_grad():
        t_h, _ = extract_features(t_sd, t_cfg, wave, lengths)
        tgt = torch.stack([t_h[i] for i in distill_layers], dim=1)
    masks = {}
    for key in gate_keys(s_sd):
        masks[key] = hc_sample(s_sd[key + ".log_alpha"], noise[key])
    s_h, _ = extract_features(s_sd, s_cfg, wave, lengths, masks)
    preds = []
    for idx, (w, b) in zip(distill_layers, heads):
        src = s_h[idx] if mode == "layer2layer" else s_h[-1]
        y = F.linear(src, w, b)
        preds.append(F.gelu(y) if mode == "predlayer" else y)
    pred = torch.stack(preds, dim=1)
    loss_d, (mse, l1, cos) = distill_loss(pred, tgt, *weights, cos_type=cos_type)
    terms = {"loss_distill": loss_d, "loss_mse": mse, "loss_l1": l1, "loss_cos": cos}
    loss = loss_d
    if lambdas is not None:
        s = 1.0 - expected_num_params(s_sd, s_cfg) / original_num_params  # lightning.py:269
        reg = lambdas[0] * (s - target_sparsity) + lambdas[1] * (s - target_sparsity) ** 2
        terms.update(loss_reg=reg, sparsity_expected=s)
        loss = loss + reg
    terms["loss"] = loss
    return loss, terms


def pruned_structure(sd, cfg):
    """Kept indices per gate after `prune()` (eval-mode masks → nonzero) — the bit-exact part of
    Wav2Vec2Model.prune model.py:115-125 / components.py:198-235, 451-483, 763-791."""
    out = {}
    for key in gate_keys(sd):
        out[key] = hc_eval_mask(sd[key + ".log_alpha"]).nonzero().squeeze(-1)
    return out
