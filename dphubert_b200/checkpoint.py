"""Checkpoint plumbing between the two training stages: what `prune.py:11-74` and
`save_final_ckpt.py:25-49` of the reference do, as functions over in-memory dicts (plus thin path wrappers).

A Lightning checkpoint of the reference stores one flat `state_dict` with the prefixes `teacher_model.`,
`student_model.`, `distill_linear_projs.` (and `lambda1` / `lambda2`); a model checkpoint is
`{"state_dict", "config", "distill_linear_projs"}`.  `DistillModule.state_dict()` of the mirror produces the same
prefixed keys, so checkpoints move between the two implementations unchanged.
"""
import copy

import torch

from .wav2vec2.model import wav2vec2_model

_GATE_FLAGS = {
    "extractor_prune_conv_channels": "feature_extractor.conv_layers.0.hard_concrete.log_alpha",
    "encoder_prune_attention_heads": "encoder.transformer.layers.0.attention.hard_concrete_for_heads.log_alpha",
    "encoder_prune_attention_layer": "encoder.transformer.layers.0.attention.hard_concrete_for_layer.log_alpha",
    "encoder_prune_feed_forward_intermediate":
        "encoder.transformer.layers.0.feed_forward.hard_concrete_for_intermediate.log_alpha",
    "encoder_prune_feed_forward_layer": "encoder.transformer.layers.0.feed_forward.hard_concrete_for_layer.log_alpha",
}


def split_prefixed(state_dict, prefix):
    """Entries of a Lightning-level state dict under `prefix`, with the prefix removed."""
    return {k[len(prefix):]: v for k, v in state_dict.items() if k.startswith(prefix)}


def pruned_config(config, prune_ret):
    """Config of the physically pruned model from `model.prune()`'s 6-tuple (prune.py:33-59): WavLM keeps
    `encoder_remaining_heads`, the others `encoder_num_heads`; all prune flags are cleared."""
    conv_config, use_attention, use_feed_forward, num_heads, remaining_heads, ff_interm_features = prune_ret
    out = copy.copy(config)
    if len(num_heads) == 0:
        assert len(remaining_heads) > 0
        out["encoder_remaining_heads"] = remaining_heads
    else:
        out["encoder_num_heads"] = num_heads
    out.update(extractor_conv_layer_config=conv_config, encoder_use_attention=use_attention,
               encoder_use_feed_forward=use_feed_forward, encoder_ff_interm_features=ff_interm_features)
    for flag in _GATE_FLAGS:
        out[flag] = False
    return out


def prune_state(lightning_state_dict, original_config):
    """prune.py:11-67 without the file I/O: rebuild the masked student (which gate families exist is read off the
    state-dict keys), load it strictly, `prune()`, and return the model checkpoint dict."""
    student_sd = split_prefixed(lightning_state_dict, "student_model.")
    config = dict(original_config)
    config.update({flag: key in student_sd for flag, key in _GATE_FLAGS.items()})
    model = wav2vec2_model(**config)
    model.load_state_dict(student_sd, strict=True)
    ret = model.prune()
    return {"state_dict": model.state_dict(), "config": pruned_config(config, ret),
            "distill_linear_projs": split_prefixed(lightning_state_dict, "distill_linear_projs.")}


def prune_from_ckpt(distilled_ckpt, original_ckpt):
    ckpt = torch.load(distilled_ckpt, map_location="cpu", weights_only=False)
    config = torch.load(original_ckpt, map_location="cpu", weights_only=False)["config"]
    return prune_state(ckpt["state_dict"], config)


def load_pruned_model(ckpt_path):
    ckpt = torch.load(ckpt_path, map_location="cpu", weights_only=False)
    model = wav2vec2_model(**ckpt["config"])
    model.load_state_dict(ckpt["state_dict"], strict=True)
    return model


def final_checkpoint(lightning_state_dict, config):
    """save_final_ckpt.py:30-46: the student after the second distillation stage with the pruned config."""
    return {"state_dict": split_prefixed(lightning_state_dict, "student_model."), "config": config,
            "distill_linear_projs": split_prefixed(lightning_state_dict, "distill_linear_projs.")}


def save_training_checkpoint(module, path=None):
    """Resumable training checkpoint in the layout of a Lightning `.ckpt` (distill.py:141-144, 180-185): the flat
    prefixed `state_dict`, `global_step`, and — under `native_step` — the Adam moments by parameter name, the optimizer
    / scheduler step counter and the noise-stream state of the fused step.  Returns the dict; writes it when `path`."""
    ckpt = {"state_dict": {k: v.detach().cpu().clone() for k, v in module.state_dict().items()},
            "global_step": int(module.global_step), "epoch": int(getattr(module, "current_epoch", 0))}
    module.on_save_checkpoint(ckpt)
    if path is not None:
        torch.save(ckpt, path)
    return ckpt


def load_training_checkpoint(module, ckpt):
    """Inverse of `save_training_checkpoint` (`ckpt`: dict or path): weights first, then the optimizer state."""
    if not isinstance(ckpt, dict):
        ckpt = torch.load(ckpt, map_location="cpu", weights_only=False)
    module.load_state_dict(ckpt["state_dict"], strict=True)
    module.on_load_checkpoint(ckpt)
    return module
