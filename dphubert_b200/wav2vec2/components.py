"""Building blocks of the prunable wav2vec2 / HuBERT / WavLM encoder — B200-native mirror of the reference
`wav2vec2/components.py`.

The classes keep the reference's names, constructor arguments, attribute paths and state-dict keys (that is the
drop-in boundary), and own the fp32 master parameters.  They do not compute with torch ops:
every `forward` hands raw device pointers to the C-ABI kernels through `dphubert_b200.engine` (channels-last
activations, fused epilogues, tcgen05 GEMMs), and raises on CPU tensors.  `get_num_params()` / `prune()` are the
reference's host-side bookkeeping (components.py:122-134,187-235,438-483,750-791,859-865,929-955,1005-1016).

Submodules are created in the same order as the reference so that a given torch seed yields the same
random initialisation.
"""
from collections import defaultdict
from typing import List, Optional, Tuple

import torch
from torch import Tensor, nn
from torch.nn import Module

from .hardconcrete import HardConcrete
from .pruning_utils import prune_conv1d_layer, prune_layer_norm, prune_linear_layer


class LayerNorm(nn.LayerNorm):
    """LayerNorm over the channel axis of a (batch, channel, time) tensor (reference components.py:54-61).
    On the native path activations are channels-last, so this is a plain row LayerNorm there."""

    def forward(self, input: Tensor) -> Tensor:
        from .. import engine

        return engine.channel_layer_norm(self, input)


class ConvLayerBlock(Module):
    """conv -> optional norm -> GELU -> optional HardConcrete channel mask (reference components.py:64-134)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride, bias, layer_norm, prune_conv_channels=False):
        super().__init__()
        self.kernel_size = kernel_size
        self.stride = stride
        self.layer_norm = layer_norm
        self.conv = nn.Conv1d(in_channels, out_channels, kernel_size, stride=stride, bias=bias)
        self.hard_concrete = HardConcrete(n_in=out_channels, init_mean=0.01) if prune_conv_channels else None

    def forward(self, x: Tensor, length: Optional[Tensor]) -> Tuple[Tensor, Optional[Tensor]]:
        """x: (batch, in_channels, in_frame) -> (batch, out_channels, out_frame), as in the reference."""
        from .. import engine

        return engine.conv_block_forward(self, x, length)

    def out_length(self, length: Optional[Tensor]) -> Optional[Tensor]:
        # integer frame arithmetic, components.py:116-119
        if length is None:
            return None
        length = torch.div(length - self.kernel_size, self.stride, rounding_mode="floor") + 1
        return torch.clamp(length, min=0)

    def get_num_params_and_out_channels(self, in_channels):
        out_channels = self.hard_concrete.l0_norm() if self.hard_concrete is not None else self.conv.out_channels
        num_params = in_channels * out_channels * self.kernel_size
        if self.conv.bias is not None:
            num_params = num_params + out_channels
        if self.layer_norm is not None:
            num_params = num_params + out_channels * 2
        return num_params, out_channels


class FeatureExtractor(Module):
    """Waveform -> frame features (reference components.py:137-235)."""

    def __init__(self, conv_layers: nn.ModuleList):
        super().__init__()
        self.conv_layers = conv_layers
        # soft mask of the last conv layer after pruning (no grad), components.py:152-156
        self.dummy_weight = nn.Parameter(
            torch.ones(conv_layers[-1].conv.out_channels, dtype=torch.float32), requires_grad=False
        )

    def forward(self, x: Tensor, length: Optional[Tensor]) -> Tuple[Tensor, Optional[Tensor]]:
        if x.ndim != 2:
            raise ValueError("Expected the input Tensor to be 2D (batch, time), " "but received {list(x.shape)}")
        from .. import engine

        return engine.feature_extractor_forward(self, x, length)

    def get_num_params_and_final_out_channels(self):
        in_channels = 1
        num_params = 0
        for layer in self.conv_layers:
            layer_params, in_channels = layer.get_num_params_and_out_channels(in_channels)
            num_params = num_params + layer_params
        num_params = num_params + in_channels  # dummy weight
        return num_params, in_channels

    def prune(self):
        """Physically drop masked conv channels (in place). Returns (conv config, kept index of the last layer)."""
        new_config = []
        index = None
        n_layers = len(self.conv_layers)
        for idx, layer in enumerate(self.conv_layers):
            hc = layer.hard_concrete
            if hc is None:
                new_config.append((layer.conv.out_channels, layer.kernel_size, layer.stride))
                index = torch.arange(layer.conv.out_channels, dtype=torch.long)
                continue
            assert not hc.training
            mask = hc()
            index = mask.nonzero().squeeze(-1)
            assert len(index) > 0, f"Conv channels pruned to zero at index {idx}"
            new_config.append((len(index), layer.kernel_size, layer.stride))
            prune_conv1d_layer(layer.conv, index, "output")
            if layer.layer_norm is not None:
                prune_layer_norm(layer.layer_norm, index)
            # the kept gates keep their soft value: fold it into the consumer of this layer's output
            if idx == n_layers - 1:
                self.dummy_weight.data *= mask
                self.dummy_weight = nn.Parameter(
                    self.dummy_weight.index_select(0, index).clone().detach(), requires_grad=False
                )
            else:
                nxt = self.conv_layers[idx + 1].conv
                nxt.weight.data *= mask.unsqueeze(-1)
                prune_conv1d_layer(nxt, index, dim="input")
            layer.hard_concrete = None
        return new_config, index


class FeatureProjection(Module):
    """LayerNorm -> Linear -> Dropout between extractor and encoder (reference components.py:238-277)."""

    def __init__(self, in_features: int, out_features: int, dropout: float):
        super().__init__()
        self.layer_norm = nn.LayerNorm(in_features)
        self.projection = nn.Linear(in_features, out_features)
        self.dropout = nn.Dropout(dropout)

    def forward(self, x):
        from .. import engine

        return engine.feature_projection_forward(self, x)

    def get_num_params(self, in_features):
        return in_features * 2 + (in_features + 1) * self.projection.out_features


class ConvolutionalPositionalEmbedding(Module):
    """Weight-normalised grouped conv positional embedding (reference components.py:280-333)."""

    def __init__(self, embed_dim: int, kernel_size: int, groups: int):
        super().__init__()
        self.embed_dim = embed_dim
        self.kernel_size = kernel_size
        conv = nn.Conv1d(embed_dim, embed_dim, kernel_size, padding=kernel_size // 2, groups=groups)
        # old-style weight norm so that the parameters are `weight_g` (1,1,k) and `weight_v`, as in the reference
        self.conv = nn.utils.weight_norm(conv, name="weight", dim=2)
        self.num_remove: int = 1 if kernel_size % 2 == 0 else 0

    def forward(self, x):
        """x: (batch, frame, feature) -> GELU(conv(x)) with the surplus last frame dropped."""
        from .. import engine

        return engine.pos_conv_forward(self, x)


def _prune_gate(layer_gate: Optional[HardConcrete], proj: nn.Linear):
    """Fold a whole-layer gate into `proj` (weight and bias). Returns True if the layer survives."""
    if layer_gate is None:
        return True
    assert not layer_gate.training
    g = layer_gate()  # (1,)
    proj.weight.data *= g
    proj.bias.data *= g
    return bool(g != 0)


class SelfAttention(Module):
    """Multi-head self attention with prunable heads / layer (reference components.py:336-483)."""

    def __init__(self, embed_dim, num_heads, head_dim, dropout=0.0, prune_heads=False, prune_layer=False):
        super().__init__()
        self.embed_dim = embed_dim
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.dropout = torch.nn.Dropout(dropout)
        self.scaling = self.head_dim**-0.5
        inner = num_heads * head_dim
        self.k_proj = nn.Linear(embed_dim, inner, bias=True)
        self.v_proj = nn.Linear(embed_dim, inner, bias=True)
        self.q_proj = nn.Linear(embed_dim, inner, bias=True)
        self.out_proj = nn.Linear(inner, embed_dim, bias=True)
        self.hard_concrete_for_heads = HardConcrete(n_in=num_heads, init_mean=0.01) if prune_heads else None
        self.hard_concrete_for_layer = HardConcrete(n_in=1, init_mean=0.01) if prune_layer else None

    def forward(self, x, attention_mask=None, position_bias=None, key_padding_mask=None):
        if x.ndim != 3 or x.shape[2] != self.embed_dim:
            raise ValueError(
                f"The expected input shape is (batch, sequence, embed_dim=={self.embed_dim}). " f"Found {x.shape}."
            )
        from .. import engine

        return engine.attention_forward(self, x, attention_mask, position_bias)

    def get_num_params(self):
        heads = self.hard_concrete_for_heads.l0_norm() if self.hard_concrete_for_heads is not None else self.num_heads
        num_params = (self.embed_dim + 1) * heads * self.head_dim * 3 + (heads * self.head_dim + 1) * self.embed_dim
        if self.hard_concrete_for_layer is not None:
            num_params = num_params * self.hard_concrete_for_layer.l0_norm()
        return num_params

    def _prune_heads(self):
        """Slice q/k/v outputs and out_proj inputs to the surviving heads. Returns the head mask."""
        head_mask = self.hard_concrete_for_heads()
        if int((head_mask != 0).sum()) > 0:
            full_mask = head_mask.repeat_interleave(self.head_dim)
            full_index = full_mask.nonzero().squeeze(-1)
            for proj in (self.k_proj, self.v_proj, self.q_proj):
                prune_linear_layer(proj, full_index, "output")
            self.out_proj.weight.data *= full_mask
            prune_linear_layer(self.out_proj, full_index, "input")
        return head_mask

    def prune(self):
        new_config = {"use_attention": True, "num_heads": self.num_heads}
        if not _prune_gate(self.hard_concrete_for_layer, self.out_proj):
            new_config["use_attention"] = False
        self.hard_concrete_for_layer = None
        if self.hard_concrete_for_heads is not None:
            assert not self.hard_concrete_for_heads.training
            head_mask = self._prune_heads()
            new_config["num_heads"] = len(head_mask.nonzero())
            if new_config["num_heads"] == 0:
                new_config["use_attention"] = False
            self.hard_concrete_for_heads = None
        return new_config


class WavLMSelfAttention(SelfAttention):
    """WavLM attention: bucketed relative position bias with a per-head GRU-style gate
    (reference components.py:486-693)."""

    def __init__(self, embed_dim, total_num_heads, remaining_heads=None, dropout=0.0, bias=True,
                 has_relative_attention_bias=False, num_buckets=32, max_distance=128, gru_rel_pos=True,
                 prune_heads=False, prune_layer=False):
        self.total_num_heads = total_num_heads
        self.remaining_heads = list(range(total_num_heads)) if remaining_heads is None else remaining_heads
        self.head_dim = embed_dim // total_num_heads
        super().__init__(embed_dim, len(self.remaining_heads), self.head_dim, dropout, prune_heads, prune_layer)
        self.has_relative_attention_bias = has_relative_attention_bias
        self.num_buckets = num_buckets
        self.max_distance = max_distance
        self.rel_attn_embed = nn.Embedding(num_buckets, total_num_heads) if has_relative_attention_bias else None
        # the reference re-creates the four projections here (to honour `bias`); same order keeps the RNG stream
        inner = len(self.remaining_heads) * self.head_dim
        self.k_proj = nn.Linear(embed_dim, inner, bias=bias)
        self.v_proj = nn.Linear(embed_dim, inner, bias=bias)
        self.q_proj = nn.Linear(embed_dim, inner, bias=bias)
        self.out_proj = nn.Linear(inner, embed_dim, bias=bias)
        self.gru_rel_pos = gru_rel_pos
        if self.gru_rel_pos:
            self.gru_rel_pos_linear = nn.Linear(self.head_dim, 8)
            self.gru_rel_pos_const = nn.Parameter(torch.ones(1, total_num_heads, 1, 1))
        self.has_position_bias = True

    def forward(self, query, attention_mask=None, position_bias=None, key_padding_mask=None):
        bsz, seq_len, embed_dim = query.size()
        assert embed_dim == self.embed_dim
        assert key_padding_mask is None
        from .. import engine

        return engine.attention_forward(self, query, attention_mask, position_bias)

    def prune(self):
        new_config = {"use_attention": True, "remaining_heads": self.remaining_heads}
        if not _prune_gate(self.hard_concrete_for_layer, self.out_proj):
            new_config["use_attention"] = False
        self.hard_concrete_for_layer = None
        if self.hard_concrete_for_heads is not None:
            assert not self.hard_concrete_for_heads.training
            head_mask = self._prune_heads()
            new_config["remaining_heads"] = head_mask.nonzero().squeeze(-1).tolist()
            if len(new_config["remaining_heads"]) == 0:
                new_config["use_attention"] = False
            self.hard_concrete_for_heads = None
        return new_config


class FeedForward(Module):
    """Linear -> GELU -> (mask) -> Linear with prunable intermediate units / layer
    (reference components.py:696-791)."""

    def __init__(self, io_features, intermediate_features, intermediate_dropout, output_dropout,
                 prune_intermediate=False, prune_layer=False):
        super().__init__()
        self.intermediate_dense = nn.Linear(io_features, intermediate_features)
        self.intermediate_dropout = nn.Dropout(intermediate_dropout)
        self.output_dense = nn.Linear(intermediate_features, io_features)
        self.output_dropout = nn.Dropout(output_dropout)
        self.hard_concrete_for_intermediate = (
            HardConcrete(n_in=intermediate_features, init_mean=0.5) if prune_intermediate else None
        )
        self.hard_concrete_for_layer = HardConcrete(n_in=1, init_mean=0.01) if prune_layer else None

    def forward(self, x):
        from .. import engine

        return engine.feed_forward_forward(self, x)

    def get_num_params(self):
        io_features = self.intermediate_dense.in_features
        if self.hard_concrete_for_intermediate is not None:
            interm = self.hard_concrete_for_intermediate.l0_norm()
        else:
            interm = self.intermediate_dense.out_features
        num_params = (io_features + 1) * interm + (interm + 1) * io_features
        if self.hard_concrete_for_layer is not None:
            num_params = num_params * self.hard_concrete_for_layer.l0_norm()
        return num_params

    def prune(self):
        new_config = {"use_feed_forward": True, "ff_interm_features": self.intermediate_dense.out_features}
        if not _prune_gate(self.hard_concrete_for_layer, self.output_dense):
            new_config["use_feed_forward"] = False
        self.hard_concrete_for_layer = None
        if self.hard_concrete_for_intermediate is not None:
            assert not self.hard_concrete_for_intermediate.training
            interm_mask = self.hard_concrete_for_intermediate()
            interm_index = interm_mask.nonzero().squeeze(-1)
            new_config["ff_interm_features"] = len(interm_index)
            if len(interm_index) == 0:
                new_config["use_feed_forward"] = False
            else:
                prune_linear_layer(self.intermediate_dense, interm_index, "output")
                self.output_dense.weight.data *= interm_mask
                prune_linear_layer(self.output_dense, interm_index, "input")
            self.hard_concrete_for_intermediate = None
        return new_config


class EncoderLayer(Module):
    """Attention + feed-forward with post-norm (Base) or pre-norm (Large) residual wiring; either sub-block may
    be None after pruning (reference components.py:794-865)."""

    def __init__(self, attention, dropout, layer_norm_first, feed_forward, embed_dim):
        super().__init__()
        self.attention = attention
        self.dropout = nn.Dropout(dropout)
        self.layer_norm = nn.LayerNorm(embed_dim)
        self.layer_norm_first = layer_norm_first
        self.feed_forward = feed_forward
        self.final_layer_norm = nn.LayerNorm(embed_dim)
        self.embed_dim = embed_dim

    def forward(self, x, attention_mask=None, position_bias=None, key_padding_mask=None):
        from .. import engine

        return engine.encoder_layer_forward(self, x, attention_mask, position_bias)

    def get_num_params(self):
        num_params = self.embed_dim * 2 * 2
        if self.attention is not None:
            num_params = num_params + self.attention.get_num_params()
        if self.feed_forward is not None:
            num_params = num_params + self.feed_forward.get_num_params()
        return num_params


class Transformer(Module):
    """Positional conv + encoder layers (reference components.py:868-955).

    NOTE: `layer_norm_first` here is the *negation* of the config flag (components.py:1283): when True (post-norm
    Base models) the LayerNorm is applied right after the positional-conv residual; otherwise only at the end of
    `forward` (never on the `get_intermediate_outputs` path)."""

    def __init__(self, pos_conv_embed, dropout, layers, layer_norm_first, layer_drop):
        super().__init__()
        self.pos_conv_embed = pos_conv_embed
        self.layer_norm = nn.LayerNorm(pos_conv_embed.embed_dim)
        self.layer_norm_first = layer_norm_first
        self.layer_drop = layer_drop
        self.dropout = nn.Dropout(dropout)
        self.layers = layers

    def _preprocess(self, x: Tensor):
        from .. import engine

        return engine.transformer_preprocess(self, x)

    def forward(self, x, attention_mask=None, position_bias=None):
        from .. import engine

        x = self._preprocess(x)
        for layer in self.layers:
            # LayerDrop: same host-side draw as the reference (components.py:902)
            if not (self.training and torch.rand(1).item() <= self.layer_drop):
                x, position_bias = layer(x, attention_mask, position_bias=position_bias)
        if not self.layer_norm_first:
            x = engine.layer_norm_module(self.layer_norm, x)
        return x

    def get_intermediate_outputs(self, x, attention_mask=None, num_layers=None, position_bias=None):
        if num_layers is not None:
            if not 0 < num_layers <= len(self.layers):
                raise ValueError(f"`num_layers` must be between [1, {len(self.layers)}]")
        ret: List[Tensor] = []
        x = self._preprocess(x)
        for layer in self.layers:
            x, position_bias = layer(x, attention_mask, position_bias=position_bias)
            ret.append(x)
            if num_layers is not None and len(ret) >= num_layers:
                return ret
        return ret

    def get_num_params(self):
        num_params = sum(p.numel() for p in self.pos_conv_embed.parameters()) + self.pos_conv_embed.embed_dim * 2
        for layer in self.layers:
            num_params = num_params + layer.get_num_params()
        return num_params

    def prune(self):
        new_config = defaultdict(list)
        for layer in self.layers:
            att = layer.attention.prune()
            new_config["use_attention"].append(att["use_attention"])
            if "remaining_heads" in att:
                new_config["remaining_heads"].append(att["remaining_heads"])
            else:
                new_config["num_heads"].append(att["num_heads"])
            if not att["use_attention"]:
                layer.attention = None
            ff = layer.feed_forward.prune()
            new_config["use_feed_forward"].append(ff["use_feed_forward"])
            new_config["ff_interm_features"].append(ff["ff_interm_features"])
            if not ff["use_feed_forward"]:
                layer.feed_forward = None
        return new_config


class Encoder(Module):
    """Feature projection + transformer (reference components.py:958-1016)."""

    def __init__(self, feature_projection, transformer):
        super().__init__()
        self.feature_projection = feature_projection
        self.transformer = transformer

    def _preprocess(self, features, lengths=None):
        from .. import engine

        return engine.encoder_preprocess(self, features, lengths)

    def forward(self, features, lengths=None):
        x, mask = self._preprocess(features, lengths)
        return self.transformer(x, attention_mask=mask)

    def extract_features(self, features, lengths=None, num_layers=None):
        x, masks = self._preprocess(features, lengths)
        interm = self.transformer.get_intermediate_outputs(x, attention_mask=masks, num_layers=num_layers)
        return [x] + interm

    def get_num_params(self, in_features):
        return self.feature_projection.get_num_params(in_features) + self.transformer.get_num_params()

    def prune(self, conv_out_index):
        prune_layer_norm(self.feature_projection.layer_norm, conv_out_index)
        prune_linear_layer(self.feature_projection.projection, conv_out_index, "input")
        return self.transformer.prune()


# ----------------------------------------------------------------------------------------------------------------
# factories (reference components.py:1020-1384)
# ----------------------------------------------------------------------------------------------------------------
def _get_feature_extractor(norm_mode, shapes, bias, prune_conv_channels=False) -> FeatureExtractor:
    if norm_mode not in ["group_norm", "layer_norm"]:
        raise ValueError("Invalid norm mode")
    blocks = []
    in_channels = 1
    for i, (out_channels, kernel_size, stride) in enumerate(shapes):
        normalization = None
        if norm_mode == "group_norm" and i == 0:
            normalization = nn.GroupNorm(num_groups=out_channels, num_channels=out_channels, affine=True)
        elif norm_mode == "layer_norm":
            normalization = LayerNorm(normalized_shape=out_channels, elementwise_affine=True)
        blocks.append(
            ConvLayerBlock(in_channels, out_channels, kernel_size, stride, bias, normalization,
                           prune_conv_channels=prune_conv_channels)
        )
        in_channels = out_channels
    return FeatureExtractor(nn.ModuleList(blocks))


def _build_encoder(make_attention, in_features, embed_dim, dropout_input, pos_conv_kernel, pos_conv_groups, num_layers,
                   use_attention, use_feed_forward, ff_interm_features, ff_interm_dropout, dropout, layer_norm_first,
                   layer_drop, prune_feed_forward_intermediate, prune_feed_forward_layer) -> Encoder:
    feature_projection = FeatureProjection(in_features, embed_dim, dropout_input)
    pos_conv = ConvolutionalPositionalEmbedding(embed_dim, pos_conv_kernel, pos_conv_groups)
    encoder_layers = nn.ModuleList()
    for idx in range(num_layers):
        attention = make_attention(idx) if use_attention[idx] else None
        feed_forward = None
        if use_feed_forward[idx]:
            feed_forward = FeedForward(embed_dim, ff_interm_features[idx], ff_interm_dropout, dropout,
                                       prune_intermediate=prune_feed_forward_intermediate,
                                       prune_layer=prune_feed_forward_layer)
        encoder_layers.append(EncoderLayer(attention, dropout, layer_norm_first, feed_forward, embed_dim))
    transformer = Transformer(pos_conv, dropout, encoder_layers, layer_norm_first=not layer_norm_first,
                              layer_drop=layer_drop)
    return Encoder(feature_projection, transformer)


def _get_encoder(in_features, embed_dim, dropout_input, pos_conv_kernel, pos_conv_groups, num_layers, use_attention,
                 use_feed_forward, num_heads, head_dim, attention_dropout, ff_interm_features, ff_interm_dropout,
                 dropout, layer_norm_first, layer_drop, prune_attention_heads=False, prune_attention_layer=False,
                 prune_feed_forward_intermediate=False, prune_feed_forward_layer=False) -> Encoder:
    def make_attention(idx):
        return SelfAttention(embed_dim, num_heads[idx], head_dim, dropout=attention_dropout,
                             prune_heads=prune_attention_heads, prune_layer=prune_attention_layer)

    return _build_encoder(make_attention, in_features, embed_dim, dropout_input, pos_conv_kernel, pos_conv_groups,
                          num_layers, use_attention, use_feed_forward, ff_interm_features, ff_interm_dropout, dropout,
                          layer_norm_first, layer_drop, prune_feed_forward_intermediate, prune_feed_forward_layer)


def _get_wavlm_encoder(in_features, embed_dim, dropout_input, pos_conv_kernel, pos_conv_groups, num_layers,
                       use_attention, use_feed_forward, total_num_heads, remaining_heads, num_buckets, max_distance,
                       attention_dropout, ff_interm_features, ff_interm_dropout, dropout, layer_norm_first, layer_drop,
                       prune_attention_heads=False, prune_attention_layer=False,
                       prune_feed_forward_intermediate=False, prune_feed_forward_layer=False) -> Encoder:
    def make_attention(i):
        return WavLMSelfAttention(embed_dim, total_num_heads[i], remaining_heads=remaining_heads[i],
                                  dropout=attention_dropout, has_relative_attention_bias=(i == 0),
                                  num_buckets=num_buckets, max_distance=max_distance,
                                  prune_heads=prune_attention_heads, prune_layer=prune_attention_layer)

    return _build_encoder(make_attention, in_features, embed_dim, dropout_input, pos_conv_kernel, pos_conv_groups,
                          num_layers, use_attention, use_feed_forward, ff_interm_features, ff_interm_dropout, dropout,
                          layer_norm_first, layer_drop, prune_feed_forward_intermediate, prune_feed_forward_layer)


def _get_padding_mask(input: Tensor, lengths: Tensor) -> Tensor:
    """(batch, max_len) bool mask, True at padded frames (reference components.py:1387-1398)."""
    batch_size, max_len, _ = input.shape
    return torch.arange(max_len, device=lengths.device).expand(batch_size, max_len) >= lengths[:, None]
