"""Ragged pruned students on the tensor-core (bf16) path: a zero-padded shadow model (dphubert_b200/padding.py)
carries the computation, the real model's parameters are views of it.  Host logic on CPU through the emulator; the
`-m gpu` twin is in test_parity_gpu.py."""
import pytest
import torch

from conftest import load_golden
from helpers import build_model, rel_err

CPU = torch.device("cpu")


@pytest.mark.parametrize("kind", ["base", "wavlm"])
def test_padded_forward_matches_reference(emulated, kind):
    g = load_golden("prune_" + kind)
    pm = build_model(g["pruned_cfg"], g["pruned_sd"], CPU, torch.bfloat16)
    with torch.no_grad():
        h, _ = pm.extract_features(g["wave"], None)
    assert len(h) == len(g["hiddens_pruned"])
    for a, b in zip(h, g["hiddens_pruned"]):
        assert a.shape == b.shape and rel_err(a, b) < 4e-2
    # the real model still has the reference's shapes and values
    sd = pm.state_dict()
    for k, v in g["pruned_sd"].items():
        assert sd[k].shape == v.shape and torch.equal(sd[k], v), k


@pytest.mark.parametrize("kind", ["base", "wavlm"])
@pytest.mark.parametrize("dtype,tol,gtol", [(torch.float32, 1e-4, 1e-3), (torch.bfloat16, 4e-2, 2e-1)])
def test_final_distill_step_on_pruned_student(emulated, kind, dtype, tol, gtol):
    """final_distill.py:61-65: distillation into the ragged pruned student; loss and gradients against the reference's
    (fp32: the ragged shapes as they are; bf16: through the padded shadow)."""
    from helpers import build_step, grad_errors, run_body

    g = load_golden("step_final_" + kind)
    step, teacher, student, projs, lam = build_step(g, CPU, dtype)
    terms = run_body(step, g, CPU)
    assert abs(terms["loss"] - float(g["loss"])) < tol * abs(float(g["loss"])), terms
    errs = grad_errors(g, student, projs, lam, floor=1e-5 if dtype == torch.float32 else 1e-3)
    assert errs and not {k: v for k, v in errs.items() if not v < gtol}, errs
    if dtype == torch.bfloat16:
        # padding stays exactly zero through the optimizer step; the real parameters moved
        from dphubert_b200 import engine

        sh = engine._SHADOWS[id(student)]
        before = {n: p.detach().clone() for n, p in student.named_parameters()}
        step._optimizer_step()
        for p, q in sh.pairs:
            if q.shape != p.shape:
                mask = torch.ones_like(q, dtype=torch.bool)
                mask[tuple(slice(0, s) for s in p.shape)] = False
                assert float(q.detach()[mask].abs().max()) == 0.0
        moved = [n for n, p in student.named_parameters() if p.requires_grad and not torch.equal(p, before[n])]
        assert len(moved) > 10
