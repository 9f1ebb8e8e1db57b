"""Generate golden fixtures from the REFERENCE itself (run in the build container, where /root/reference exists):

    python oracle/make_golden.py            # writes tests/golden/*.pt

For small seeded configurations of each model family it runs the reference's own modules
(`/root/reference/wav2vec2`, `/root/reference/lightning.py` behind a 6-line pytorch_lightning stub) and stores
inputs + the reference's outputs: per-layer hidden states, the `_step` loss terms, gradients of selected tensors,
`get_num_params()`, and the `prune()` result.  It also asserts that `oracle/dphubert_oracle.py` reproduces every
stored value, which is what pins the oracle.  The fixtures are consumed by tests/ (CPU: oracle vs golden; GPU: CUDA
path vs golden) without the reference being present.
"""
import copy
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("DPH_REFERENCE", "/root/reference")
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, HERE)

import dphubert_oracle as O  # noqa: E402


def import_reference():
    sys.path.insert(0, REF)
    # lightning.py imports pytorch_lightning (not installed) and the dataset module (needs torchaudio I/O)
    pl = types.ModuleType("pytorch_lightning")

    class LightningModule(torch.nn.Module):
        global_step = 0
        current_epoch = 0

        def log_dict(self, *a, **k):
            pass

    pl.LightningModule = LightningModule
    sys.modules["pytorch_lightning"] = pl
    ds = types.ModuleType("dataset.audio_dataset")
    for n in ("BucketizeBatchSampler", "DistributedBatchSampler", "CollateFnAudio", "AudioDataset"):
        setattr(ds, n, object)
    sys.modules["dataset"] = types.ModuleType("dataset")
    sys.modules["dataset.audio_dataset"] = ds
    import importlib.util

    from wav2vec2.model import wav2vec2_model  # noqa

    spec = importlib.util.spec_from_file_location("ref_lightning", os.path.join(REF, "lightning.py"))
    rl = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rl)
    return wav2vec2_model, rl


def small_cfg(kind):
    """Scaled-down configs with the exact structure of the canonical ones (convert_hubert_from_hf.py:18-44,
    convert_hubert_large_from_fairseq.py:20-46, convert_wavlm_from_hf.py:19-47)."""
    L, D, H, Dh, Fi, C = 2, 192, 3, 64, 128, 64  # head_dim 64 so the bf16 runs take the fused attention kernel
    conv = [(C, 10, 5)] + [(C, 3, 2)] * 4 + [(C, 2, 2)] * 2
    cfg = dict(
        extractor_mode="group_norm", extractor_conv_layer_config=conv, extractor_conv_bias=False,
        encoder_embed_dim=D, encoder_projection_dropout=0.0, encoder_pos_conv_kernel=16, encoder_pos_conv_groups=4,
        encoder_num_layers=L, encoder_use_attention=[True] * L, encoder_use_feed_forward=[True] * L,
        encoder_num_heads=[H] * L, encoder_head_dim=Dh, encoder_attention_dropout=0.0,
        encoder_ff_interm_features=[Fi] * L, encoder_ff_interm_dropout=0.0, encoder_dropout=0.0,
        encoder_layer_norm_first=False, encoder_layer_drop=0.0, aux_num_out=None, normalize_waveform=False,
        extractor_prune_conv_channels=False, encoder_prune_attention_heads=False,
        encoder_prune_attention_layer=False, encoder_prune_feed_forward_intermediate=False,
        encoder_prune_feed_forward_layer=False,
    )
    if kind == "large":
        cfg.update(extractor_mode="layer_norm", encoder_layer_norm_first=True, normalize_waveform=True,
                   extractor_conv_bias=True)
    if kind == "wavlm":
        del cfg["encoder_num_heads"], cfg["encoder_head_dim"]
        cfg.update(encoder_total_num_heads=[H] * L, encoder_remaining_heads=[list(range(H)) for _ in range(L)],
                   encoder_num_buckets=32, encoder_max_distance=80)
    return cfg


def randomize(model, seed):
    """non-trivial values for every parameter (norm affine params and biases are not left at 1 / 0)"""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if "log_alpha" in n or n.endswith("dummy_weight"):
                continue
            if p.dim() == 1 or "gru_rel_pos_const" in n:
                p.add_(0.1 * torch.randn(p.shape, generator=g))
            else:
                p.mul_(1.5)


def student_cfg(cfg, units=("conv", "head", "interm")):
    c = copy.deepcopy(cfg)
    c.update(extractor_prune_conv_channels="conv" in units, encoder_prune_attention_heads="head" in units,
             encoder_prune_attention_layer="attlayer" in units,
             encoder_prune_feed_forward_intermediate="interm" in units,
             encoder_prune_feed_forward_layer="ffnlayer" in units)
    return c


def sd_of(model):
    return {k: v.detach().clone() for k, v in model.state_dict().items()}


def close(a, b, tol=2e-5, what="", floor=1e-12):
    a, b = torch.as_tensor(a).float(), torch.as_tensor(b).float()
    err = (a - b).abs().max().item() / max(b.abs().max().item(), floor)
    assert err < tol, "%s: oracle deviates from the reference by %.3e" % (what, err)
    return err


def sample_grad(g):
    g = g.detach().flatten()
    stride = max(1, g.numel() // 4096)
    return g[::stride][:4096].clone()


def case_forward(wav2vec2_model, kind, seed):
    cfg = small_cfg(kind)
    torch.manual_seed(seed)
    model = wav2vec2_model(**cfg).eval()
    randomize(model, seed + 1)
    g = torch.Generator().manual_seed(2022)
    wave = 0.1 * torch.randn(2, 8000, generator=g)
    lengths = torch.tensor([8000, 6500])
    sd = sd_of(model)
    out = {"cfg": cfg, "state_dict": sd, "wave": wave, "lengths": lengths}
    with torch.no_grad():
        h1, l1 = model.extract_features(wave, lengths)
        h2, _ = model.extract_features(wave, None)
        y, _ = model(wave, lengths)
    out.update(hiddens_len=[h.clone() for h in h1], out_lengths=l1, hiddens_nolen=[h.clone() for h in h2],
               forward_out=y.clone(), num_params=int(model.get_num_params()),
               numel=sum(p.numel() for p in model.parameters()))
    # pin the oracle
    with torch.no_grad():
        o1, ol = O.extract_features(sd, cfg, wave, lengths)
        o2, _ = O.extract_features(sd, cfg, wave, None)
    assert torch.equal(ol, l1)
    e = max(close(a, b, what="%s hidden(len)" % kind) for a, b in zip(o1, h1))
    e = max(e, max(close(a, b, what="%s hidden" % kind) for a, b in zip(o2, h2)))
    assert int(O.expected_num_params(sd, cfg)) == out["num_params"]
    print("forward %-6s: %d hiddens of %s, oracle max rel err %.2e, params %d" %
          (kind, len(h1), tuple(h1[0].shape), e, out["num_params"]))
    return out


def case_step(wav2vec2_model, rl, kind, seed, units=("conv", "head", "interm"), use_reg=True, mode="layer2layer",
              pruned_from=None):
    """One `_step` of the reference. `pruned_from` = a prune_*.pt fixture: the student is that ragged pruned model
    (final_distill.py:61-65), no gates, use_reg False."""
    cfg = small_cfg(kind)
    torch.manual_seed(seed)
    teacher = wav2vec2_model(**cfg).eval()
    randomize(teacher, seed + 1)
    for p in teacher.parameters():
        p.requires_grad = False
    if pruned_from is not None:
        assert not use_reg
        scfg = copy.deepcopy(pruned_from["pruned_cfg"])
        student = wav2vec2_model(**scfg)
        student.load_state_dict(pruned_from["pruned_sd"], strict=True)
    else:
        scfg = student_cfg(cfg, units if use_reg else ())
        student = wav2vec2_model(**scfg)
        student.load_state_dict(teacher.state_dict(), strict=False)
    g = torch.Generator().manual_seed(seed + 2)
    with torch.no_grad():
        for n, p in student.named_parameters():
            if "log_alpha" in n:  # spread the gates so that the clamp's both sides and the interior are exercised
                p.copy_(torch.randn(p.shape, generator=g) * 2.0 + 1.0)
            elif p.requires_grad:
                p.add_(0.02 * torch.randn(p.shape, generator=g))
    student.train()
    L = cfg["encoder_num_layers"]
    D = cfg["encoder_embed_dim"]
    layers = [0, 1, L]
    shared = torch.nn.Linear(D, D)
    first = torch.nn.Linear(D, D)
    with torch.no_grad():
        for lin in (shared, first):
            lin.weight.copy_(torch.eye(D) + 0.05 * torch.randn(D, D, generator=g))
            lin.bias.copy_(0.05 * torch.randn(D, generator=g))
    if mode == "layer2layer":
        projs = torch.nn.ModuleList([first, shared, shared])
    else:
        projs = torch.nn.ModuleList([torch.nn.Sequential(copy.deepcopy(first), torch.nn.GELU()) for _ in layers])
    loss_mod = rl.DistillLoss(l2_weight=0.5, l1_weight=1.0, cos_weight=1.0, cos_type="raw" if kind != "wavlm" else "log_sig")
    module = rl.DistillModule(
        teacher_model=teacher, student_model=student, distill_mode=mode, distill_layers=layers,
        distill_linear_projs=projs, distill_loss=loss_mod, learning_rate=2e-4, weight_decay=0.0, warmup_updates=10,
        max_updates=100, use_reg=use_reg, reg_learning_rate=0.02, target_sparsity=0.75, sparsity_warmup_updates=0,
        tsv_dir=".", train_subset="x", seconds_per_batch=1, num_workers=0)
    if use_reg:
        with torch.no_grad():
            module.lambda1.fill_(0.3)
            module.lambda2.fill_(0.2)
    wave = 0.1 * torch.randn(2, 8000, generator=torch.Generator().manual_seed(7))
    lengths = torch.tensor([8000, 8000])
    # identical HardConcrete noise on both sides: pre-draw u per gate and feed it through Tensor.uniform_
    gates = [n[: -len(".log_alpha")] for n, _ in student.named_parameters() if n.endswith(".log_alpha")]
    noise = {k: torch.empty(student.get_parameter(k + ".log_alpha").shape).uniform_(1e-6, 1 - 1e-6, generator=g)
             for k in gates}
    mods = dict(student.named_modules())
    for k in gates:
        hc = mods[k]

        def fwd(hc=hc, u=noise[k]):
            s = torch.sigmoid((torch.log(u / (1 - u)) + hc.log_alpha) / hc.beta)
            return (s * (hc.limit_r - hc.limit_l) + hc.limit_l).clamp(min=0.0, max=1.0)

        hc.forward = fwd  # same arithmetic as hardconcrete.py:96-99 with the noise injected
    s_sd0 = sd_of(student)
    loss = module._step((wave, lengths), 0, "train")
    loss.backward()
    grads = {"student_model." + n: sample_grad(p.grad) for n, p in student.named_parameters() if p.grad is not None}
    seen = {}
    for i, pr in enumerate(projs):
        for n, p in pr.named_parameters():
            if id(p) not in seen:
                seen[id(p)] = True
                grads["distill_linear_projs.%d.%s" % (i, n)] = sample_grad(p.grad)
    if use_reg:
        grads["lambda1"] = module.lambda1.grad.clone()
        grads["lambda2"] = module.lambda2.grad.clone()
    heads = []
    for pr in projs:
        lin = pr if isinstance(pr, torch.nn.Linear) else pr[0]
        heads.append((lin.weight.detach().clone(), lin.bias.detach().clone()))
    head_ids = []
    uniq = []
    for pr in projs:
        lin = pr if isinstance(pr, torch.nn.Linear) else pr[0]
        if id(lin) not in uniq:
            uniq.append(id(lin))
        head_ids.append(uniq.index(id(lin)))
    out = {"cfg": cfg, "student_cfg": scfg, "teacher_sd": sd_of(teacher), "student_sd": s_sd0, "wave": wave,
           "lengths": lengths, "noise": noise, "layers": layers, "heads": heads, "head_ids": head_ids, "mode": mode,
           "weights": (0.5, 1.0, 1.0), "cos_type": loss_mod.cos_type, "use_reg": use_reg,
           "lambdas": (0.3, 0.2) if use_reg else None, "target_sparsity": 0.75,
           "original_num_params": module.original_num_params, "loss": loss.detach().clone(), "grads": grads}
    # pin the oracle: same loss and same gradients from the functional restatement
    s_sd = {k: v.clone().requires_grad_(v.is_floating_point() and "dummy" not in k) for k, v in s_sd0.items()}
    uniq_heads = {}
    o_heads = []
    for hid, (w, b) in zip(head_ids, heads):
        if hid not in uniq_heads:
            uniq_heads[hid] = (w.clone().requires_grad_(True), b.clone().requires_grad_(True))
        o_heads.append(uniq_heads[hid])
    lam = [torch.tensor(0.3, requires_grad=True), torch.tensor(0.2, requires_grad=True)] if use_reg else None
    o_loss, terms = O.distill_step_loss(out["teacher_sd"], cfg, s_sd, scfg, o_heads, layers, wave, lengths, noise, lam,
                                        0.75, module.original_num_params, out["weights"], out["cos_type"], mode)
    o_loss.backward()
    e = close(o_loss, loss, what="%s step loss" % kind)
    ge = 0.0
    for n, p in student.named_parameters():
        if p.grad is not None:
            ge = max(ge, close(sample_grad(s_sd[n].grad), grads["student_model." + n], tol=2e-4, what="grad " + n,
                               floor=1e-5))  # k_proj.bias has an analytically zero gradient (softmax shift invariance)
    if use_reg:
        close(lam[0].grad, grads["lambda1"], what="lambda1 grad")
        close(lam[1].grad, grads["lambda2"], what="lambda2 grad")
    out["terms"] = {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in terms.items()}
    print("step    %-6s reg=%s %s: loss %.6f, oracle rel err loss %.2e grads %.2e, %d gates" %
          (kind, use_reg, mode, loss.item(), e, ge, len(gates)))
    return out


def case_prune(wav2vec2_model, kind, seed, units=("conv", "head", "interm", "attlayer", "ffnlayer")):
    cfg = small_cfg(kind)
    scfg = student_cfg(cfg, units)
    torch.manual_seed(seed)
    model = wav2vec2_model(**scfg)
    randomize(model, seed + 1)
    g = torch.Generator().manual_seed(seed + 2)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if "log_alpha" in n:
                if p.numel() == 1:  # whole-layer gates: close one attention layer and one FFN layer
                    closed = (".layers.1.attention." in n) or (".layers.0.feed_forward." in n)
                    p.fill_(-6.0 if closed else 4.0)
                else:
                    p.copy_(torch.randn(p.shape, generator=g) * 3.0 + 0.5)
    sd0 = sd_of(model)
    wave = 0.1 * torch.randn(2, 8000, generator=torch.Generator().manual_seed(11))
    model.eval()
    with torch.no_grad():
        h_masked, _ = model.extract_features(wave, None)  # eval-mode (compiled) masks, un-pruned
    ret = model.prune()
    conv_config, use_att, use_ff, num_heads, remaining, ff_dims = ret
    pruned_sd = sd_of(model)
    # the reference re-builds the pruned model from the new config (prune.py:33-74) before using it
    pcfg = copy.deepcopy(scfg)
    if len(num_heads) == 0:
        pcfg["encoder_remaining_heads"] = remaining
    else:
        pcfg["encoder_num_heads"] = num_heads
    pcfg.update(extractor_conv_layer_config=conv_config, encoder_use_attention=use_att,
                encoder_use_feed_forward=use_ff, encoder_ff_interm_features=ff_dims,
                extractor_prune_conv_channels=False, encoder_prune_attention_heads=False,
                encoder_prune_attention_layer=False, encoder_prune_feed_forward_intermediate=False,
                encoder_prune_feed_forward_layer=False)
    pmodel = wav2vec2_model(**pcfg)
    pmodel.load_state_dict(pruned_sd, strict=True)
    pmodel.eval()
    with torch.no_grad():
        h_pruned, _ = pmodel.extract_features(wave, None)
        o_pruned, _ = O.extract_features(pruned_sd, pcfg, wave, None)
    for a, b in zip(o_pruned, h_pruned):
        close(a, b, what="pruned-model hidden")
    kept = O.pruned_structure(sd0, scfg)
    # pin the oracle's kept indices on the reference's result
    for i, (c, k, s) in enumerate(conv_config):
        idx = kept["feature_extractor.conv_layers.%d.hard_concrete" % i]
        assert len(idx) == c, (i, len(idx), c)
    w0 = sd0["feature_extractor.conv_layers.0.conv.weight"]
    assert torch.equal(pruned_sd["feature_extractor.conv_layers.0.conv.weight"],
                       w0.index_select(0, kept["feature_extractor.conv_layers.0.hard_concrete"]))
    for l in range(cfg["encoder_num_layers"]):
        pre = "encoder.transformer.layers.%d." % l
        if use_ff[l]:
            idx = kept[pre + "feed_forward.hard_concrete_for_intermediate"]
            assert len(idx) == ff_dims[l]
            assert torch.equal(pruned_sd[pre + "feed_forward.intermediate_dense.bias"],
                               sd0[pre + "feed_forward.intermediate_dense.bias"].index_select(0, idx))
        if use_att[l]:
            idx = kept[pre + "attention.hard_concrete_for_heads"]
            if kind == "wavlm":
                assert idx.tolist() == remaining[l]
            else:
                assert len(idx) == num_heads[l]
    out = {"student_cfg": scfg, "state_dict": sd0, "wave": wave, "prune_ret": [list(map(tuple, conv_config)),
           list(use_att), list(use_ff), list(num_heads), [list(r) for r in remaining], list(ff_dims)],
           "pruned_sd": pruned_sd, "kept": kept, "hiddens_masked": [h.clone() for h in h_masked],
           "hiddens_pruned": [h.clone() for h in h_pruned], "kind": kind, "pruned_cfg": pcfg}
    print("prune   %-6s: conv %s heads %s%s ff %s att %s ffn %s" %
          (kind, [c for c, _, _ in conv_config], num_heads, remaining if kind == "wavlm" else "", ff_dims, use_att, use_ff))
    return out


def case_aux_layerdrop(wav2vec2_model, seed=400):
    """Public `forward` (model.py:127-169) with the `aux` head and train-mode LayerDrop (components.py:894-907): the
    reference draws `torch.rand(1).item()` once per layer from the global CPU generator, so seeding it pins the keep
    pattern on both sides."""
    cfg = small_cfg("base")
    L = 4
    cfg.update(encoder_num_layers=L, encoder_use_attention=[True] * L, encoder_use_feed_forward=[True] * L,
               encoder_num_heads=[3] * L, encoder_ff_interm_features=[128] * L, aux_num_out=29,
               encoder_layer_drop=0.5)
    torch.manual_seed(seed)
    model = wav2vec2_model(**cfg)
    randomize(model, seed + 1)
    wave = 0.1 * torch.randn(2, 8000, generator=torch.Generator().manual_seed(seed + 2))
    lengths = torch.tensor([8000, 7000])
    out = {"cfg": cfg, "state_dict": sd_of(model), "wave": wave, "lengths": lengths, "runs": []}
    model.eval()
    with torch.no_grad():
        y, ol = model(wave, lengths)
    out["eval_out"], out["out_lengths"] = y.clone(), ol.clone()
    model.train()
    for s in (1, 2, 3, 4):
        torch.manual_seed(1000 + s)
        with torch.no_grad():
            y, _ = model(wave, lengths)
        torch.manual_seed(1000 + s)
        keep = [not (torch.rand(1).item() <= 0.5) for _ in range(L)]
        out["runs"].append({"seed": 1000 + s, "keep": keep, "out": y.clone()})
    assert len({tuple(r["keep"]) for r in out["runs"]}) > 1
    print("aux/layerdrop: out %s, keep patterns %s" % (tuple(y.shape), [r["keep"] for r in out["runs"]]))
    return out


def final_distill_cases(wav2vec2_model, rl):
    """Second training stage: the pruned (ragged) students of the prune_* fixtures distilled from a full teacher."""
    for kind, seed in (("base", 300), ("wavlm", 320)):
        pruned = torch.load(os.path.join(OUT, "prune_%s.pt" % kind), weights_only=False)
        torch.save(case_step(wav2vec2_model, rl, kind, seed, use_reg=False, pruned_from=pruned),
                   os.path.join(OUT, "step_final_%s.pt" % kind))


def sampler_cases():
    """Batch lists / collated batches of the reference's dataset/audio_dataset.py for seeded inputs."""
    sys.path.insert(0, REF)
    from dataset.audio_dataset import BucketizeBatchSampler, CollateFnAudio, DistributedBatchSampler

    g = torch.Generator().manual_seed(41)
    lengths = torch.randint(20000, 260000, (700,), generator=g).tolist()
    out = {"lengths": lengths, "cases": []}
    for kw, seed in ((dict(num_buckets=50, min_len=32000, max_len=250000, max_token_count=16000 * 40, shuffle=True), 5),
                     (dict(num_buckets=1000, min_len=32000, max_len=250000, max_token_count=16000 * 160, shuffle=True), 6),
                     (dict(num_buckets=10, batch_size=7, shuffle=False, drop_last=True), 7),
                     (dict(num_buckets=13, batch_size=16, shuffle=True, drop_last=False), 8)):
        torch.manual_seed(seed)
        bs = BucketizeBatchSampler(lengths, **kw)
        case = {"kw": kw, "seed": seed, "iter_list": [list(b) for b in bs.iter_list], "len": bs.__len__(), "dist": []}
        for dkw in (dict(num_replicas=4, shuffle=True, seed=3, drop_last=False),
                    dict(num_replicas=3, shuffle=True, seed=1, drop_last=True),
                    dict(num_replicas=2, shuffle=False, drop_last=False)):
            subsets = []
            for rank in range(dkw["num_replicas"]):
                torch.manual_seed(seed)
                fresh = BucketizeBatchSampler(lengths, **kw)  # shuffle=False extends the sampler's list in place
                ds = DistributedBatchSampler(fresh, rank=rank, **dkw)
                subsets.append([list(b) for b in ds])
            case["dist"].append({"kw": dkw, "subsets": subsets})
        out["cases"].append(case)
    # collate: small utterances so the fixture stays small
    wl = torch.randint(700, 3000, (48,), generator=g).tolist()
    waves = [0.1 * torch.randn(1, n, generator=g) for n in wl]
    torch.manual_seed(9)
    bs = BucketizeBatchSampler(wl, num_buckets=6, max_token_count=12000, shuffle=True)
    batches = [list(b) for b in bs.iter_list]
    coll = []
    for pad, rand_crop, seed in ((False, True, 21), (False, False, 22), (True, True, 23)):
        torch.manual_seed(seed)
        fn = CollateFnAudio(pad=pad, rand_crop=rand_crop)
        res = [fn([(waves[i], wl[i]) for i in b]) for b in batches]
        coll.append({"pad": pad, "rand_crop": rand_crop, "seed": seed, "waveforms": [r[0].clone() for r in res],
                     "lengths": [r[1].clone() for r in res]})
    out["collate"] = {"wave_lengths": wl, "waves": [w.clone() for w in waves], "batches": batches, "runs": coll}
    torch.save(out, os.path.join(OUT, "samplers.pt"))
    print("samplers: %d sampler cases, %d collate batches" % (len(out["cases"]), len(batches)))


def main():
    if "--samplers-only" in sys.argv:
        os.makedirs(OUT, exist_ok=True)
        return sampler_cases()
    wav2vec2_model, rl = import_reference()
    os.makedirs(OUT, exist_ok=True)
    if "--aux-only" in sys.argv:
        return torch.save(case_aux_layerdrop(wav2vec2_model), os.path.join(OUT, "forward_base_aux.pt"))
    if "--final-only" in sys.argv:  # adds the final-distill fixtures without rewriting the others
        return final_distill_cases(wav2vec2_model, rl)
    for kind, seed in (("base", 0), ("large", 10), ("wavlm", 20)):
        torch.save(case_forward(wav2vec2_model, kind, seed), os.path.join(OUT, "forward_%s.pt" % kind))
    for kind, seed in (("base", 100), ("large", 110), ("wavlm", 120)):
        torch.save(case_step(wav2vec2_model, rl, kind, seed), os.path.join(OUT, "step_%s.pt" % kind))
    torch.save(case_step(wav2vec2_model, rl, "base", 130, use_reg=False), os.path.join(OUT, "step_base_noreg.pt"))
    torch.save(case_step(wav2vec2_model, rl, "base", 140, units=("conv", "head", "interm", "attlayer", "ffnlayer"),
                         mode="predlayer"), os.path.join(OUT, "step_base_allunits_predlayer.pt"))
    for kind, seed in (("base", 200), ("wavlm", 220)):
        torch.save(case_prune(wav2vec2_model, kind, seed), os.path.join(OUT, "prune_%s.pt" % kind))
    final_distill_cases(wav2vec2_model, rl)
    torch.save(case_aux_layerdrop(wav2vec2_model), os.path.join(OUT, "forward_base_aux.pt"))
    sampler_cases()
    # known-answer anchors of the full-size models: parameter counts
    from wav2vec2.model import wav2vec2_model as ref_build
    sys.path.insert(0, ROOT)
    from dphubert_b200.wav2vec2.model import hubert_base_config, hubert_large_config, wavlm_base_plus_config
    anchors = {}
    for name, c in (("hubert_base", hubert_base_config()), ("hubert_large", hubert_large_config()),
                    ("wavlm_base_plus", wavlm_base_plus_config())):
        m = ref_build(**c)
        anchors[name] = {"numel": sum(p.numel() for p in m.parameters()), "get_num_params": int(m.get_num_params()),
                         "keys": sorted(m.state_dict().keys())}
        print("anchor", name, anchors[name]["numel"], anchors[name]["get_num_params"])
    torch.save(anchors, os.path.join(OUT, "anchors.pt"))


if __name__ == "__main__":
    main()
