"""Golden fixtures at BASELINE.json's REAL configurations, generated from the REFERENCE itself:

    python oracle/make_golden_real.py         # writes tests/golden/real_*.pt   (build container only)

The small fixtures of `make_golden.py` pin every code path on a 2-layer toy model; this script pins the *sizes*
BASELINE.json names: HuBERT Base 12L/768d (config 1: 4 x 4 s fp32 forward; a 10 s forward so that
T = 499 runs the multi-tile attention / tensor-core positional conv / 512-channel conv-0 paths inside the model),
the Base prune step at real dimensions (config 3 on 2 x 4 s), WavLM Base+ with the real 320-bucket / 800-distance
table (config 4: forward at T = 499 and a prune step) and HuBERT Large 24L/1024d (config 5: forward + prune step).

The weights are NOT stored (95-315 M parameters): they are the torch default initialisation under a stated
`torch.manual_seed`, which the mirror package reproduces bit for bit (same constructors in the same order); the
fixture carries a float64 checksum per tensor so the test can prove it holds the same weights.  Stored are the
reference's outputs, strided sub-samples of every hidden state / gradient plus full-tensor norms.  The script also
asserts that `oracle/dphubert_oracle.py` reproduces the reference at these sizes.
"""
import copy
import os
import sys
import time

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import dphubert_oracle as O  # noqa: E402
from make_golden import close, import_reference, sd_of, student_cfg  # noqa: E402

NS = 16384   # samples kept per hidden state
NG = 4096    # samples kept per gradient tensor


def sample(t, n):
    t = t.detach().flatten()
    s = max(1, t.numel() // n)
    return t[::s][:n].clone()


def checksum(sd):
    """{key: (sum, sum of |x|)} in float64 — proves two state dicts hold the same values without storing them."""
    return {k: (float(v.double().sum()), float(v.double().abs().sum())) for k, v in sd.items()
            if torch.is_tensor(v) and v.is_floating_point()}


def stats(t):
    t = t.detach().double()
    return {"l2": float(t.norm()), "max": float(t.abs().max()), "numel": t.numel()}


def make_wave(B, T0, seed):
    return 0.1 * torch.randn(B, T0, generator=torch.Generator().manual_seed(seed))


def configs():
    from dphubert_b200.wav2vec2.model import hubert_base_config, hubert_large_config, wavlm_base_plus_config

    return {"base": hubert_base_config(), "large": hubert_large_config(), "wavlm": wavlm_base_plus_config()}


def case_forward(build, kind, seed, B, T0, wave_seed, lengths):
    cfg = configs()[kind]
    torch.manual_seed(seed)
    model = build(**cfg).eval()
    sd = sd_of(model)
    wave = make_wave(B, T0, wave_seed)
    ln = None if lengths is None else torch.tensor(lengths)
    t0 = time.time()
    with torch.no_grad():
        h, ol = model.extract_features(wave, ln)
    t_ref = time.time() - t0
    with torch.no_grad():
        oh, ool = O.extract_features(sd, cfg, wave, ln)
    e = max(close(a, b, what="%s hidden %d" % (kind, i)) for i, (a, b) in enumerate(zip(oh, h)))
    if ln is not None:
        assert torch.equal(ol, ool)
    out = {"kind": kind, "seed": seed, "B": B, "T0": T0, "wave_seed": wave_seed, "lengths": lengths,
           "wave_sum": float(wave.double().sum()), "weights": checksum(sd), "out_lengths": ol,
           "shape": tuple(h[0].shape), "hiddens": [sample(x, NS) for x in h], "stats": [stats(x) for x in h],
           "oracle_vs_reference": e}
    print("forward %-6s %dx%d: %d hiddens %s, reference %.1f s, oracle rel err %.2e" %
          (kind, B, T0, len(h), tuple(h[0].shape), t_ref, e))
    return out


def case_step(build, rl, kind, seed, B, T0, wave_seed, layers, groups, gtol=5e-4, perturb=0.0, spread=False):
    """One reference `_step` (lightning.py:245-296) at real dimensions: student = teacher weights + freshly
    initialised gates (distill.py:69-84), identity-initialised shared heads per layer group (distill.py:24-26,
    93-99), l1 = 1, cos = 1 (raw), l2 = 0, target sparsity 0.75 reached (run.sh:21-35)."""
    cfg = configs()[kind]
    torch.manual_seed(seed)
    teacher = build(**cfg).eval()
    for p in teacher.parameters():
        p.requires_grad = False
    scfg = student_cfg(cfg, ("conv", "head", "interm"))
    torch.manual_seed(seed + 1)
    student = build(**scfg)
    res = student.load_state_dict(teacher.state_dict(), strict=False)
    assert all("log_alpha" in k for k in res.missing_keys) and not res.unexpected_keys
    if perturb:
        # Large only: with student == teacher and identity heads most |pred - target| of the pre-norm 24-layer model
        # sit at fp32 round-off, where the L1 term's sign() is noise (two fp32 evaluation orders then disagree by
        # 1e-2 in the late layers' gradients); a seeded weight perturbation moves the comparison off that edge
        pg = torch.Generator().manual_seed(seed + 3)
        with torch.no_grad():
            for n, p in student.named_parameters():
                if "log_alpha" not in n and p.requires_grad:
                    p.add_(perturb * torch.randn(p.shape, generator=pg))
    if spread:  # gates spread so that both clamps of the HardConcrete stretch and its interior are exercised
        sg = torch.Generator().manual_seed(seed + 4)
        with torch.no_grad():
            for n, p in student.named_parameters():
                if "log_alpha" in n:
                    p.copy_(torch.randn(p.shape, generator=sg) * 2.0 + 1.0)
    student.train()
    for m in student.modules():  # parity configuration: dropout off (SURVEY.md §8c "noise control")
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
        if hasattr(m, "layer_drop"):
            m.layer_drop = 0.0
    D = cfg["encoder_embed_dim"]
    projs = torch.nn.ModuleList()
    head_ids = []
    for gi, g in enumerate(groups):
        lin = torch.nn.Linear(D, D)
        with torch.no_grad():
            lin.weight.copy_(torch.eye(D))
            lin.bias.fill_(0)
        for _ in g:
            projs.append(lin)
            head_ids.append(gi)
    loss_mod = rl.DistillLoss(l2_weight=0.0, l1_weight=1.0, cos_weight=1.0, cos_type="raw")
    module = rl.DistillModule(
        teacher_model=teacher, student_model=student, distill_mode="layer2layer", distill_layers=layers,
        distill_linear_projs=projs, distill_loss=loss_mod, learning_rate=2e-4, weight_decay=0.0, warmup_updates=15000,
        max_updates=50000, use_reg=True, reg_learning_rate=0.02, target_sparsity=0.75, sparsity_warmup_updates=0,
        tsv_dir=".", train_subset="x", seconds_per_batch=1, num_workers=0)
    with torch.no_grad():
        module.lambda1.fill_(0.3)
        module.lambda2.fill_(0.2)
    wave = make_wave(B, T0, wave_seed)
    lengths = torch.tensor([T0] * B)
    g = torch.Generator().manual_seed(seed + 2)
    gates = [n[: -len(".log_alpha")] for n, _ in student.named_parameters() if n.endswith(".log_alpha")]
    noise = {k: torch.empty(student.get_parameter(k + ".log_alpha").shape).uniform_(1e-6, 1 - 1e-6, generator=g)
             for k in gates}
    mods = dict(student.named_modules())
    for k in gates:
        hc = mods[k]

        def fwd(hc=hc, u=noise[k]):
            s = torch.sigmoid((torch.log(u / (1 - u)) + hc.log_alpha) / hc.beta)
            return (s * (hc.limit_r - hc.limit_l) + hc.limit_l).clamp(min=0.0, max=1.0)

        hc.forward = fwd  # hardconcrete.py:96-99 with the noise injected
    s_sd0 = sd_of(student)
    t0 = time.time()
    loss = module._step((wave, lengths), 0, "train")
    loss.backward()
    t_ref = time.time() - t0
    grads, gstats = {}, {}
    for n, p in student.named_parameters():
        if p.grad is not None:
            grads["student_model." + n] = sample(p.grad, NG)
            gstats["student_model." + n] = stats(p.grad)
    seen = set()
    for i, pr in enumerate(projs):
        for n, p in pr.named_parameters():
            if id(p) not in seen:
                seen.add(id(p))
                grads["distill_linear_projs.%d.%s" % (i, n)] = sample(p.grad, NG)
                gstats["distill_linear_projs.%d.%s" % (i, n)] = stats(p.grad)
    grads["lambda1"] = module.lambda1.grad.clone()
    grads["lambda2"] = module.lambda2.grad.clone()
    # ---- pin the oracle at this size
    t_sd = sd_of(teacher)
    o_sd = {k: v.clone().requires_grad_(v.is_floating_point() and "dummy" not in k) for k, v in s_sd0.items()}
    uniq = {}
    o_heads = []
    for hid in head_ids:
        if hid not in uniq:
            uniq[hid] = (torch.eye(D).requires_grad_(True), torch.zeros(D, requires_grad=True))
        o_heads.append(uniq[hid])
    lam = [torch.tensor(0.3, requires_grad=True), torch.tensor(0.2, requires_grad=True)]
    o_loss, terms = O.distill_step_loss(t_sd, cfg, o_sd, scfg, o_heads, layers, wave, lengths, noise, lam, 0.75,
                                        module.original_num_params, (0.0, 1.0, 1.0), "raw", "layer2layer")
    o_loss.backward()
    e = close(o_loss, loss, what="%s step loss" % kind)
    ge = 0.0
    for n, p in student.named_parameters():
        if p.grad is not None:
            ge = max(ge, close(sample(o_sd[n].grad, NG), grads["student_model." + n], tol=gtol, what="grad " + n,
                               floor=1e-6))
    close(lam[0].grad, grads["lambda1"], what="lambda1 grad")
    close(lam[1].grad, grads["lambda2"], what="lambda2 grad")
    out = {"kind": kind, "seed": seed, "B": B, "T0": T0, "wave_seed": wave_seed, "wave_sum": float(wave.double().sum()),
           "layers": layers, "head_ids": head_ids, "perturb": perturb, "spread": spread, "noise": noise, "lambdas": (0.3, 0.2), "target_sparsity": 0.75,
           "original_num_params": module.original_num_params, "teacher_weights": checksum(t_sd),
           "student_weights": checksum(s_sd0), "loss": loss.detach().clone(),
           "terms": {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in terms.items()},
           "grads": grads, "grad_stats": gstats, "oracle_vs_reference": {"loss": e, "grads": ge}}
    print("step    %-6s %dx%d: loss %.6f sparsity %.5f, reference %.1f s, oracle rel err loss %.2e grads %.2e, %d gates"
          % (kind, B, T0, loss.item(), float(terms["sparsity_expected"]), t_ref, e, ge, len(gates)))
    return out


def main():
    build, rl = import_reference()
    os.makedirs(OUT, exist_ok=True)
    only = [a for a in sys.argv[1:] if not a.startswith("-")]

    def want(name):
        return not only or name in only

    def save(name, obj):
        torch.save(obj, os.path.join(OUT, name + ".pt"))

    if want("real_base_fwd"):      # BASELINE config 1 exactly
        save("real_base_fwd", case_forward(build, "base", 0, 4, 64000, 2022, [64000] * 4))
    if want("real_base_fwd10"):    # T = 499, ragged lengths
        save("real_base_fwd10", case_forward(build, "base", 0, 2, 160000, 2023, [160000, 131000]))
    if want("real_base_step"):     # config 3 at the CPU arm's sample
        save("real_base_step", case_step(build, rl, "base", 0, 2, 64000, 2024, [0, 4, 8, 12], [[0], [4, 8, 12]]))
    if want("real_base_step_p"):   # same, well-conditioned: perturbed student weights, spread gates
        save("real_base_step_p", case_step(build, rl, "base", 0, 2, 64000, 2024, [0, 4, 8, 12], [[0], [4, 8, 12]],
                                           perturb=0.02, spread=True))
    if want("real_wavlm_fwd10"):   # config 4: real bucket table at T = 499
        save("real_wavlm_fwd10", case_forward(build, "wavlm", 0, 2, 160000, 2025, None))
    if want("real_wavlm_step"):
        save("real_wavlm_step", case_step(build, rl, "wavlm", 0, 2, 64000, 2026, [0, 4, 8, 12], [[0], [4, 8, 12]],
                                          perturb=0.02, spread=True))
    if want("real_large_fwd"):     # config 5 model
        save("real_large_fwd", case_forward(build, "large", 0, 2, 64000, 2027, [64000, 50000]))
    if want("real_large_step"):
        save("real_large_step", case_step(build, rl, "large", 0, 1, 64000, 2028, [0, 4, 8, 12, 16, 20, 24],
                                          [[0], [4, 8, 12, 16, 20, 24]], gtol=2e-3, perturb=0.02, spread=True))


if __name__ == "__main__":
    main()
