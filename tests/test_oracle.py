"""The oracle (oracle/dphubert_oracle.py) against the fixtures produced by the reference itself
(oracle/make_golden.py). Runs without the reference and without a GPU."""
import pytest
import torch

import dphubert_oracle as O
from conftest import load_golden
from helpers import rel_err, sample_grad

TOL = 2e-5  # fp32 CPU, same torch build on both sides


@pytest.mark.parametrize("kind", ["base", "large", "wavlm"])
def test_forward_matches_reference(kind):
    g = load_golden("forward_" + kind)
    with torch.no_grad():
        h, l = O.extract_features(g["state_dict"], g["cfg"], g["wave"], g["lengths"])
        h2, _ = O.extract_features(g["state_dict"], g["cfg"], g["wave"], None)
    assert torch.equal(l, g["out_lengths"])
    assert len(h) == len(g["hiddens_len"]) == g["cfg"]["encoder_num_layers"] + 1
    for a, b in zip(h, g["hiddens_len"]):
        assert rel_err(a, b) < TOL
    for a, b in zip(h2, g["hiddens_nolen"]):
        assert rel_err(a, b) < TOL
    assert int(O.expected_num_params(g["state_dict"], g["cfg"])) == g["num_params"]


@pytest.mark.parametrize("name", ["step_base", "step_large", "step_wavlm", "step_base_noreg",
                                  "step_base_allunits_predlayer"])
def test_step_loss_and_grads_match_reference(name):
    g = load_golden(name)
    s_sd = {k: v.clone().requires_grad_(v.is_floating_point() and "dummy" not in k) for k, v in g["student_sd"].items()}
    uniq, heads = {}, []
    for hid, (w, b) in zip(g["head_ids"], g["heads"]):
        if hid not in uniq:
            uniq[hid] = (w.clone().requires_grad_(True), b.clone().requires_grad_(True))
        heads.append(uniq[hid])
    lam = None
    if g["use_reg"]:
        lam = [torch.tensor(g["lambdas"][0], requires_grad=True), torch.tensor(g["lambdas"][1], requires_grad=True)]
    loss, terms = O.distill_step_loss(g["teacher_sd"], g["cfg"], s_sd, g["student_cfg"], heads, g["layers"], g["wave"],
                                      g["lengths"], g["noise"], lam, g["target_sparsity"], g["original_num_params"],
                                      g["weights"], g["cos_type"], g["mode"])
    loss.backward()
    assert rel_err(loss, g["loss"]) < TOL
    for k, ref in g["grads"].items():
        if k.startswith("student_model."):
            got = s_sd[k[len("student_model."):]].grad
            assert rel_err(sample_grad(got), ref, floor=1e-5) < 2e-4, k
    if lam is not None:
        assert rel_err(lam[0].grad, g["grads"]["lambda1"]) < TOL
        assert rel_err(lam[1].grad, g["grads"]["lambda2"]) < TOL


@pytest.mark.parametrize("kind", ["base", "wavlm"])
def test_pruned_structure_is_bit_exact(kind):
    g = load_golden("prune_" + kind)
    kept = O.pruned_structure(g["state_dict"], g["student_cfg"])
    for k, idx in g["kept"].items():
        assert torch.equal(kept[k], idx), k
    conv_config, use_att, use_ff, num_heads, remaining, ff_dims = g["prune_ret"]
    for i, (c, _, _) in enumerate(conv_config):
        assert len(kept["feature_extractor.conv_layers.%d.hard_concrete" % i]) == c
    with torch.no_grad():
        h, _ = O.extract_features(g["pruned_sd"], g["pruned_cfg"], g["wave"], None)
    for a, b in zip(h, g["hiddens_pruned"]):
        assert rel_err(a, b) < TOL


def test_hardconcrete_known_answers():
    # constants of hardconcrete.py:28-36,58-64
    assert abs(O.HC_BIAS - (2 / 3) * 2.3978952727983707) < 1e-12
    la = torch.tensor([-20.0, 0.0, 20.0])
    u = torch.tensor([0.5, 0.5, 0.5])
    m = O.hc_sample(la, u)
    assert m[0] == 0.0 and m[2] == 1.0 and abs(m[1].item() - 0.5) < 1e-6
    # banker's rounding of the expected number of zeros (hardconcrete.py:105-106)
    la = torch.full((4,), -O.HC_BIAS)  # sigmoid(0) = .5 each -> 2 expected open -> round(2.0) = 2 zeros
    assert int((O.hc_eval_mask(la) == 0).sum()) == 2
