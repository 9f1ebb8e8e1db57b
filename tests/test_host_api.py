"""Host-side drop-in surface (no GPU): config → module tree, state-dict keys, parameter counts, prune()."""
import pytest
import torch

from conftest import load_golden
from dphubert_b200.wav2vec2 import model as M


@pytest.mark.parametrize("name,cfg", [("hubert_base", M.hubert_base_config), ("hubert_large", M.hubert_large_config),
                                      ("wavlm_base_plus", M.wavlm_base_plus_config)])
def test_state_dict_keys_and_counts_match_reference(name, cfg):
    anchors = load_golden("anchors")[name]
    m = M.wav2vec2_model(**cfg())
    assert sorted(m.state_dict().keys()) == anchors["keys"]
    assert sum(p.numel() for p in m.parameters()) == anchors["numel"]
    assert int(m.get_num_params()) == anchors["get_num_params"]


def test_known_answer_param_counts():
    # 94,371,456 (Base) / 315,434,624 (Large); WavLM numel 94,381,680 but get_num_params 94,371,456
    a = load_golden("anchors")
    assert a["hubert_base"]["numel"] == 94371456 and a["hubert_large"]["numel"] == 315434624
    assert a["wavlm_base_plus"]["numel"] == 94381680 and a["wavlm_base_plus"]["get_num_params"] == 94371456


def test_masked_student_has_31_gate_modules():
    cfg = M.hubert_base_config()
    cfg.update(extractor_prune_conv_channels=True, encoder_prune_attention_heads=True,
               encoder_prune_feed_forward_intermediate=True)
    m = M.wav2vec2_model(**cfg)
    gates = [(n, p) for n, p in m.named_parameters() if "log_alpha" in n]
    assert len(gates) == 31 and sum(p.numel() for _, p in gates) == 40592


@pytest.mark.parametrize("kind", ["base", "wavlm"])
def test_prune_matches_reference_bit_exact(kind):
    g = load_golden("prune_" + kind)
    m = M.wav2vec2_model(**g["student_cfg"])
    m.load_state_dict(g["state_dict"], strict=True)
    conv_config, use_att, use_ff, num_heads, remaining, ff_dims = m.prune()
    ref = g["prune_ret"]
    assert [tuple(c) for c in conv_config] == [tuple(c) for c in ref[0]]
    assert list(use_att) == ref[1] and list(use_ff) == ref[2]
    assert list(num_heads) == ref[3] and [list(r) for r in remaining] == ref[4] and list(ff_dims) == ref[5]
    sd = m.state_dict()
    assert sorted(sd.keys()) == sorted(g["pruned_sd"].keys())
    for k, v in g["pruned_sd"].items():
        assert torch.equal(sd[k], v), k
    # strict reload into a model built from the pruned config (prune.py:70-74)
    pm = M.wav2vec2_model(**g["pruned_cfg"])
    pm.load_state_dict(sd, strict=True)


def test_cpu_tensors_are_refused():
    from dphubert_b200 import _lib

    g = load_golden("forward_base")
    m = M.wav2vec2_model(**g["cfg"])
    with pytest.raises(_lib.DphError):
        m.extract_features(g["wave"], None)


@pytest.mark.parametrize("kind", ["base", "wavlm"])
def test_prune_checkpoint_flow_matches_reference(kind, tmp_path):
    """prune.py:11-74 / save_final_ckpt.py:25-49 as functions: same pruned config, same tensors, and the result
    loads back strictly."""
    from dphubert_b200 import checkpoint as C

    g = load_golden("prune_" + kind)
    base_cfg = {k: v for k, v in g["student_cfg"].items()}
    for flag in C._GATE_FLAGS:  # the original (teacher) config has no gates; prune_state re-derives them from the keys
        base_cfg[flag] = False
    lsd = {"student_model." + k: v for k, v in g["state_dict"].items()}
    lsd.update({"distill_linear_projs.0.weight": torch.eye(3), "lambda1": torch.tensor(0.0)})
    out = C.prune_state(lsd, base_cfg)
    ref_cfg = g["pruned_cfg"]
    assert set(out["config"]) == set(ref_cfg)
    for k, v in ref_cfg.items():
        got = out["config"][k]
        if k == "extractor_conv_layer_config":
            assert [tuple(c) for c in got] == [tuple(c) for c in v]
        else:
            assert got == v, k
    assert sorted(out["state_dict"]) == sorted(g["pruned_sd"])
    for k, v in g["pruned_sd"].items():
        assert torch.equal(out["state_dict"][k], v), k
    assert list(out["distill_linear_projs"]) == ["0.weight"]
    path = tmp_path / "pruned.pth"
    torch.save(out, path)
    m = C.load_pruned_model(path)
    fin = C.final_checkpoint({"student_model." + k: v for k, v in m.state_dict().items()}, out["config"])
    assert sorted(fin["state_dict"]) == sorted(g["pruned_sd"])
