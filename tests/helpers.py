"""Shared test plumbing: build mirror models / native steps from golden fixtures and compare against them."""
import torch

from dphubert_b200.trainer import NativeDistillStep
from dphubert_b200.wav2vec2.model import wav2vec2_model


def rel_err(got, ref, floor=1e-12):
    got, ref = torch.as_tensor(got).detach().float().cpu(), torch.as_tensor(ref).detach().float().cpu()
    return (got - ref).abs().max().item() / max(ref.abs().max().item(), floor)


def sample_grad(t):
    """same sub-sampling as oracle/make_golden.py:sample_grad"""
    t = t.detach().flatten()
    s = max(1, t.numel() // 4096)
    return t[::s][:4096]


def build_model(cfg, sd, device, dtype, train=False, frozen=False):
    m = wav2vec2_model(**cfg)
    res = m.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    m = m.to(device)
    m.set_compute_dtype(dtype)
    m.train(train)
    if frozen:
        for p in m.parameters():
            p.requires_grad = False
    return m


def build_step(g, device, dtype, **kw):
    """teacher / student / heads / NativeDistillStep from a `step_*.pt` fixture, HardConcrete noise injected."""
    teacher = build_model(g["cfg"], g["teacher_sd"], device, dtype, train=False, frozen=True)
    student = build_model(g["student_cfg"], g["student_sd"], device, dtype, train=True)
    D = g["cfg"]["encoder_embed_dim"]
    uniq, projs = {}, []
    for hid, (w, b) in zip(g["head_ids"], g["heads"]):
        if hid not in uniq:
            lin = torch.nn.Linear(D, D)
            with torch.no_grad():
                lin.weight.copy_(w)
                lin.bias.copy_(b)
            uniq[hid] = lin.to(device)
        lin = uniq[hid]
        projs.append(lin if g["mode"] == "layer2layer" else torch.nn.Sequential(lin, torch.nn.GELU()))
    lam = None
    if g["use_reg"]:
        lam = [torch.nn.Parameter(torch.tensor(g["lambdas"][0])), torch.nn.Parameter(torch.tensor(g["lambdas"][1]))]
    w2, w1, wc = g["weights"]
    args = dict(target_sparsity=g["target_sparsity"], sparsity_warmup_updates=0, warmup_updates=10, max_updates=100)
    args.update(kw)
    step = NativeDistillStep(teacher, student, g["layers"], projs, g["mode"], w2, w1, wc, g["cos_type"], g["use_reg"], lam,
                             g["original_num_params"], **args)
    inject_noise(student, g["noise"])
    return step, teacher, student, projs, lam


def inject_noise(student, noise):
    mods = dict(student.named_modules())
    for k, u in noise.items():
        mods[k].noise = u


def run_body(step, g, device):
    """forward + backward of one step (no optimizer), returns the loss terms"""
    step._push_scalars()
    step._body(g["wave"].to(device), g["lengths"].to(device))
    return step.loss_terms()


def grad_errors(g, student, projs, lam, floor=1e-5):
    """{name: relative error} of every gradient stored in the fixture"""
    errs = {}
    for n, p in student.named_parameters():
        key = "student_model." + n
        if key in g["grads"]:
            errs[key] = rel_err(sample_grad(p.grad), g["grads"][key], floor)
    seen = set()
    for i, pr in enumerate(projs):
        lin = pr if isinstance(pr, torch.nn.Linear) else pr[0]
        if id(lin) in seen:
            continue
        seen.add(id(lin))
        for nm, p in (("weight", lin.weight), ("bias", lin.bias)):
            for key in ("distill_linear_projs.%d.%s" % (i, nm), "distill_linear_projs.%d.0.%s" % (i, nm)):
                if key in g["grads"]:
                    errs[key] = rel_err(sample_grad(p.grad), g["grads"][key], floor)
    if lam is not None:
        errs["lambda1"] = rel_err(lam[0].grad, g["grads"]["lambda1"])
        errs["lambda2"] = rel_err(lam[1].grad, g["grads"]["lambda2"])
    return errs
