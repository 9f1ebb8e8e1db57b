"""`Wav2Vec2Model` and its factories — B200-native mirror of the reference `wav2vec2/model.py`.

Drop-in surface: `wav2vec2_model(**config)` (dispatching to `wavlm_model` when the config has
`encoder_remaining_heads`, model.py:172-178), `load_state_dict`, `extract_features` returning the per-layer hidden
states, `forward`, `get_num_params()` (differentiable expected size) and `prune()`.

`extract_features` runs the whole-model native path (`dphubert_b200.engine`): one autograd node per block whose
forward/backward are sequences of C-ABI kernel launches.  Extension over the reference: `compute_dtype`
(torch.float32 = fp32 parity mode on CUDA cores, torch.bfloat16 = tcgen05 tensor-core mode).
"""
from typing import List, Optional, Tuple

import torch
from torch import Tensor
from torch.nn import Module

from . import components


class Wav2Vec2Model(Module):
    def __init__(self, normalize_waveform: bool, feature_extractor: Module, encoder: Module,
                 aux: Optional[Module] = None):
        super().__init__()
        self.normalize_waveform = normalize_waveform
        self.feature_extractor = feature_extractor
        self.encoder = encoder
        self.aux = aux
        self.compute_dtype = torch.float32

    def set_compute_dtype(self, dtype):
        """torch.float32 (parity mode) or torch.bfloat16 (tensor-core mode)."""
        assert dtype in (torch.float32, torch.bfloat16)
        self.compute_dtype = dtype
        return self

    def extract_features(self, waveforms: Tensor, lengths: Optional[Tensor] = None,
                         num_layers: Optional[int] = None) -> Tuple[List[Tensor], Optional[Tensor]]:
        """(batch, samples) waveforms -> ([hidden_0 .. hidden_L], frame lengths)  (reference model.py:57-107)."""
        from .. import engine

        return engine.model_extract_features(self, waveforms, lengths, num_layers)

    def get_num_params(self):
        """Current (expected, differentiable when HardConcrete gates are present) parameter count
        (reference model.py:109-113)."""
        from .. import engine

        fused = engine.fused_num_params(self)
        if fused is not None:
            return fused
        feature_extractor_size, encoder_in_features = self.feature_extractor.get_num_params_and_final_out_channels()
        return feature_extractor_size + self.encoder.get_num_params(encoder_in_features)

    def prune(self):
        """Turn the learned masks into a physically smaller model, in place (reference model.py:115-125)."""
        self.eval()
        conv_config, conv_out_index = self.feature_extractor.prune()
        transformer_config = self.encoder.prune(conv_out_index)
        from .. import engine

        engine.invalidate(self)
        return (
            conv_config,
            transformer_config["use_attention"],
            transformer_config["use_feed_forward"],
            transformer_config["num_heads"],
            transformer_config["remaining_heads"],
            transformer_config["ff_interm_features"],
        )

    def forward(self, waveforms: Tensor, lengths: Optional[Tensor] = None) -> Tuple[Tensor, Optional[Tensor]]:
        """Encoder output (after LayerDrop / trailing LayerNorm / aux head), reference model.py:127-169."""
        from .. import engine

        return engine.model_forward(self, waveforms, lengths)


def wav2vec2_model(**configs) -> Wav2Vec2Model:
    """Build from a checkpoint config dict; WavLM is selected by the presence of `encoder_remaining_heads`."""
    if "encoder_remaining_heads" in configs:
        return wavlm_model(**configs)
    return wav2vec2_model_original(**configs)


_DEFAULT_CONV = [(512, 10, 5)] + [(512, 3, 2)] * 4 + [(512, 2, 2)] * 2


def wav2vec2_model_original(
    extractor_mode: str,
    extractor_conv_layer_config: Optional[List[Tuple[int, int, int]]],
    extractor_conv_bias: bool,
    encoder_embed_dim: int,
    encoder_projection_dropout: float,
    encoder_pos_conv_kernel: int,
    encoder_pos_conv_groups: int,
    encoder_num_layers: int,
    encoder_use_attention: List[bool],
    encoder_use_feed_forward: List[bool],
    encoder_num_heads: List[int],
    encoder_head_dim: int,
    encoder_attention_dropout: float,
    encoder_ff_interm_features: List[int],
    encoder_ff_interm_dropout: float,
    encoder_dropout: float,
    encoder_layer_norm_first: bool,
    encoder_layer_drop: float,
    aux_num_out: Optional[int],
    normalize_waveform: bool,
    extractor_prune_conv_channels: bool = False,
    encoder_prune_attention_heads: bool = False,
    encoder_prune_attention_layer: bool = False,
    encoder_prune_feed_forward_intermediate: bool = False,
    encoder_prune_feed_forward_layer: bool = False,
) -> Wav2Vec2Model:
    """wav2vec 2.0 / HuBERT builder — same keyword arguments as the reference (model.py:181-361)."""
    if extractor_conv_layer_config is None:
        extractor_conv_layer_config = list(_DEFAULT_CONV)
    feature_extractor = components._get_feature_extractor(
        extractor_mode, extractor_conv_layer_config, extractor_conv_bias,
        prune_conv_channels=extractor_prune_conv_channels,
    )
    encoder = components._get_encoder(
        in_features=extractor_conv_layer_config[-1][0],
        embed_dim=encoder_embed_dim,
        dropout_input=encoder_projection_dropout,
        pos_conv_kernel=encoder_pos_conv_kernel,
        pos_conv_groups=encoder_pos_conv_groups,
        num_layers=encoder_num_layers,
        use_attention=encoder_use_attention,
        use_feed_forward=encoder_use_feed_forward,
        num_heads=encoder_num_heads,
        head_dim=encoder_head_dim,
        attention_dropout=encoder_attention_dropout,
        ff_interm_features=encoder_ff_interm_features,
        ff_interm_dropout=encoder_ff_interm_dropout,
        dropout=encoder_dropout,
        layer_norm_first=encoder_layer_norm_first,
        layer_drop=encoder_layer_drop,
        prune_attention_heads=encoder_prune_attention_heads,
        prune_attention_layer=encoder_prune_attention_layer,
        prune_feed_forward_intermediate=encoder_prune_feed_forward_intermediate,
        prune_feed_forward_layer=encoder_prune_feed_forward_layer,
    )
    aux = None
    if aux_num_out is not None:
        aux = torch.nn.Linear(in_features=encoder_embed_dim, out_features=aux_num_out)
    return Wav2Vec2Model(normalize_waveform, feature_extractor, encoder, aux)


def wavlm_model(
    extractor_mode: str,
    extractor_conv_layer_config: Optional[List[Tuple[int, int, int]]],
    extractor_conv_bias: bool,
    encoder_embed_dim: int,
    encoder_projection_dropout: float,
    encoder_pos_conv_kernel: int,
    encoder_pos_conv_groups: int,
    encoder_num_layers: int,
    encoder_use_attention: List[bool],
    encoder_use_feed_forward: List[bool],
    encoder_total_num_heads: List[int],
    encoder_remaining_heads: List[List[int]],
    encoder_num_buckets: int,
    encoder_max_distance: int,
    encoder_attention_dropout: float,
    encoder_ff_interm_features: List[int],
    encoder_ff_interm_dropout: float,
    encoder_dropout: float,
    encoder_layer_norm_first: bool,
    encoder_layer_drop: float,
    aux_num_out: Optional[int],
    normalize_waveform: bool,
    extractor_prune_conv_channels: bool = False,
    encoder_prune_attention_heads: bool = False,
    encoder_prune_attention_layer: bool = False,
    encoder_prune_feed_forward_intermediate: bool = False,
    encoder_prune_feed_forward_layer: bool = False,
) -> Wav2Vec2Model:
    """WavLM builder — same keyword arguments as the reference (model.py:736-862)."""
    if extractor_conv_layer_config is None:
        extractor_conv_layer_config = list(_DEFAULT_CONV)
    feature_extractor = components._get_feature_extractor(
        extractor_mode, extractor_conv_layer_config, extractor_conv_bias,
        prune_conv_channels=extractor_prune_conv_channels,
    )
    encoder = components._get_wavlm_encoder(
        in_features=extractor_conv_layer_config[-1][0],
        embed_dim=encoder_embed_dim,
        dropout_input=encoder_projection_dropout,
        pos_conv_kernel=encoder_pos_conv_kernel,
        pos_conv_groups=encoder_pos_conv_groups,
        num_layers=encoder_num_layers,
        use_attention=encoder_use_attention,
        use_feed_forward=encoder_use_feed_forward,
        total_num_heads=encoder_total_num_heads,
        remaining_heads=encoder_remaining_heads,
        num_buckets=encoder_num_buckets,
        max_distance=encoder_max_distance,
        attention_dropout=encoder_attention_dropout,
        ff_interm_features=encoder_ff_interm_features,
        ff_interm_dropout=encoder_ff_interm_dropout,
        dropout=encoder_dropout,
        layer_norm_first=encoder_layer_norm_first,
        layer_drop=encoder_layer_drop,
        prune_attention_heads=encoder_prune_attention_heads,
        prune_attention_layer=encoder_prune_attention_layer,
        prune_feed_forward_intermediate=encoder_prune_feed_forward_intermediate,
        prune_feed_forward_layer=encoder_prune_feed_forward_layer,
    )
    aux = None
    if aux_num_out is not None:
        aux = torch.nn.Linear(in_features=encoder_embed_dim, out_features=aux_num_out)
    return Wav2Vec2Model(normalize_waveform, feature_extractor, encoder, aux)


# ----------------------------------------------------------------------------------------------------------------
# canonical config dicts (the reference keeps them inside its converters:
# convert_hubert_from_hf.py:18-44, convert_hubert_large_from_fairseq.py:20-46, convert_wavlm_from_hf.py:19-47)
# ----------------------------------------------------------------------------------------------------------------
def hubert_base_config() -> dict:
    return dict(
        extractor_mode="group_norm", extractor_conv_layer_config=list(_DEFAULT_CONV), extractor_conv_bias=False,
        encoder_embed_dim=768, encoder_projection_dropout=0.1, encoder_pos_conv_kernel=128,
        encoder_pos_conv_groups=16, encoder_num_layers=12, encoder_use_attention=[True] * 12,
        encoder_use_feed_forward=[True] * 12, encoder_num_heads=[12] * 12, encoder_head_dim=64,
        encoder_attention_dropout=0.1, encoder_ff_interm_features=[3072] * 12, encoder_ff_interm_dropout=0.0,
        encoder_dropout=0.1, encoder_layer_norm_first=False, encoder_layer_drop=0.05, aux_num_out=None,
        normalize_waveform=False, extractor_prune_conv_channels=False, encoder_prune_attention_heads=False,
        encoder_prune_attention_layer=False, encoder_prune_feed_forward_intermediate=False,
        encoder_prune_feed_forward_layer=False,
    )


def hubert_large_config() -> dict:
    cfg = hubert_base_config()
    cfg.update(
        extractor_mode="layer_norm", encoder_embed_dim=1024, encoder_num_layers=24,
        encoder_use_attention=[True] * 24, encoder_use_feed_forward=[True] * 24, encoder_num_heads=[16] * 24,
        encoder_ff_interm_features=[4096] * 24, encoder_layer_norm_first=True, normalize_waveform=True,
    )
    return cfg


def wavlm_base_plus_config() -> dict:
    cfg = hubert_base_config()
    del cfg["encoder_num_heads"], cfg["encoder_head_dim"]
    cfg.update(
        encoder_total_num_heads=[12] * 12, encoder_remaining_heads=[list(range(12)) for _ in range(12)],
        encoder_num_buckets=320, encoder_max_distance=800,
    )
    return cfg


def hubert_base(**overrides) -> Wav2Vec2Model:
    """HuBERT Base preset (reference model.py:538-596)."""
    cfg = hubert_base_config()
    cfg.update(overrides)
    return wav2vec2_model(**cfg)


def hubert_large(**overrides) -> Wav2Vec2Model:
    cfg = hubert_large_config()
    cfg.update(overrides)
    return wav2vec2_model(**cfg)


def wavlm_base(**overrides) -> Wav2Vec2Model:
    cfg = wavlm_base_plus_config()
    cfg.update(overrides)
    return wav2vec2_model(**cfg)
