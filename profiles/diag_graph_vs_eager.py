"""Diagnostic: per-parameter gradient difference between the eager launch sequence and captured-graph replays of the
SAME forward + backward (`NativeDistillStep._body`) over interleaved batch shapes on one shared workspace.

    python profiles/diag_graph_vs_eager.py [--no-streams] [--repeat N]

Prints, per step, the loss difference and the parameters whose gradients differ most (relative to the largest gradient
entry of the model), plus the replay-to-replay spread of the graph twin.  Used to tell real cross-stream races from
atomics-order round-off when `tests/test_trainer_gpu.py::test_shape_keyed_graph_cache_matches_eager` trips.

What it found (profiles/r02_diag_l1_sign.txt): the forward pass is reproducible to fp32 round-off only, and the L1 term's
gradient w1 * sign(pred - target) / N jumps when an element has pred - target ~ 0: at step 7 of the test's sequence one
such element makes the gradient norm take one of three equally spaced values (3.585123 / 3.585318 / 3.585513) from run
to run — in the eager twin as often as in the graph twin, with side streams, PDL or extra synchronisation on or off.
With `--smooth` (MSE + cosine) every step agrees to 1e-7."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from conftest import load_golden  # noqa: E402
from helpers import build_step  # noqa: E402
from dphubert_b200 import ops  # noqa: E402

DEV = torch.device("cuda:0")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-streams", action="store_true")
    ap.add_argument("--repeat", type=int, default=3)
    ap.add_argument("--fixture", default="step_base")
    ap.add_argument("--full", action="store_true", help="the whole step through run_step / step() (optimizer included); "
                    "gradients are snapshotted right before the optimizer launches")
    ap.add_argument("--sync-before", default="", help="comma list of dphubert_b200.ops functions to synchronise before")
    ap.add_argument("--sync-after", default="", help="comma list of dphubert_b200.ops functions to synchronise after")
    ap.add_argument("--smooth", action="store_true", help="MSE + cosine only (no L1 term, whose sign() jumps at 0)")
    ap.add_argument("--spy", action="store_true", help="--first-replay: copy the loss backward's inputs before it runs")
    ap.add_argument("--first-replay", action="store_true", help="replay a freshly captured graph twice on the same "
                    "input and noise epoch and list every workspace buffer that differs between the two replays")
    args = ap.parse_args()
    g = load_golden(args.fixture)
    from dphubert_b200 import gemm as G

    def wrap(name, before):
        mod = G if name.startswith("gemm") else ops
        fn = getattr(mod, name.split(".")[-1])

        def w(*a, **k):
            if before and not torch.cuda.is_current_stream_capturing():
                torch.cuda.synchronize()
            r = fn(*a, **k)
            if not before and not torch.cuda.is_current_stream_capturing():
                torch.cuda.synchronize()
            return r

        setattr(mod, name.split(".")[-1], w)

    for nm in filter(None, args.sync_before.split(",")):
        wrap(nm, True)
    for nm in filter(None, args.sync_after.split(",")):
        wrap(nm, False)
    if args.smooth:
        from dphubert_b200 import trainer as TR
        init = TR.NativeDistillStep.__init__

        def patched(self, *a, **k):
            init(self, *a, **k)
            self.w2, self.w1 = 1.0, 0.0

        TR.NativeDistillStep.__init__ = patched
    if args.first_replay:
        return first_replay(args, g)
    if args.full:
        return full(args, g)

    def fresh():
        ops.manual_seed(21)
        ops.rng_set(0)
        s = build_step(g, DEV, torch.float32, learning_rate=0.0, reg_learning_rate=0.0)
        for h in s[2].modules():
            if hasattr(h, "noise"):
                h.noise = None
        if args.no_streams:
            s[0].two_streams = False
        return s

    shapes = [8000, 5000, 8000, 5000, 6500, 8000, 5000, 6500, 8000, 6500, 5000, 6500, 8000, 5000] * args.repeat

    def batch(i, T0):
        return ((g["wave"][:, :T0] * (1.0 + 0.01 * (i % 11))).contiguous().to(DEV), torch.tensor([T0, T0 - 300]).to(DEV))

    a, b = fresh(), fresh()
    names = [n for n, _ in a[2].named_parameters()]
    graphs, seen = {}, {}
    allocs = (b[0].At, b[0].As, b[0].Ah)
    worst_all = 0.0
    for i, T0 in enumerate(shapes):
        wave, ln = batch(i, T0)
        for s in (a, b):
            s[0].zero_grad()
            s[0]._push_scalars()
        ops.rng_set(100 + i)
        a[0]._body(wave, ln)
        torch.cuda.synchronize()
        ops.rng_set(100 + i)
        mode = "replay"
        if T0 not in graphs:
            n = seen.get(T0, 0)
            seen[T0] = n + 1
            for al in allocs:
                al.always_zero = True
            if n == 0:
                mode = "eager"
                b[0]._body(wave, ln)
            else:
                mode = "capture+replay"
                w_in = b[0].As("in_wave", tuple(wave.shape), torch.float32)
                l_in = b[0].As("in_lengths", (2,), torch.int64)
                torch.cuda.synchronize()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr):
                    b[0]._body(w_in, l_in)
                graphs[T0] = (gr, w_in, l_in)
            for al in allocs:
                al.always_zero = False
        if mode != "eager":
            gr, w_in, l_in = graphs[T0]
            w_in.copy_(wave)
            l_in.copy_(ln)
            gr.replay()
        torch.cuda.synchronize()
        la, lb = a[0].loss_terms()["loss"], b[0].loss_terms()["loss"]
        gmax = max(float(p.grad.abs().max()) for p in a[2].parameters() if p.grad is not None)
        diffs = []
        for n, p, q in zip(names, a[2].parameters(), b[2].parameters()):
            if p.grad is None:
                continue
            diffs.append((float((p.grad - q.grad).abs().max()) / gmax, n))
        diffs.sort(reverse=True)
        na = float(torch.sqrt(sum((p.grad.double() ** 2).sum() for p in a[2].parameters() if p.grad is not None)))
        nb = float(torch.sqrt(sum((p.grad.double() ** 2).sum() for p in b[2].parameters() if p.grad is not None)))
        worst_all = max(worst_all, diffs[0][0])
        flag = " <<<" if diffs[0][0] > 1e-5 or abs(na - nb) > 1e-5 * na else ""
        print("step %2d T0=%d %-14s dloss=%.2e dgnorm=%.2e worst=%.2e %s%s" %
              (i, T0, mode, abs(la - lb) / abs(la), abs(na - nb) / na, diffs[0][0], diffs[0][1], flag))
        if flag:
            for d, n in diffs[:8]:
                print("        %.3e  %s" % (d, n))
    print("worst over the run: %.3e" % worst_all)


def full(args, g):
    def fresh():
        ops.manual_seed(21)
        ops.rng_set(0)
        s = build_step(g, DEV, torch.float32, learning_rate=0.0, reg_learning_rate=0.0)
        for h in s[2].modules():
            if hasattr(h, "noise"):
                h.noise = None
        if args.no_streams:
            s[0].two_streams = False
        st = s[0]
        snap = [torch.zeros_like(t) for t in st._grad_buffers()]
        orig = st._optimizer_step

        def spy():
            for d, t in zip(snap, st._grad_buffers()):
                d.copy_(t)
            orig()

        st._optimizer_step = spy
        return s, snap

    shapes = [8000, 5000, 8000, 5000, 6500, 8000, 5000, 6500, 8000, 6500, 5000] * args.repeat

    def batch(i, T0):
        return ((g["wave"][:, :T0] * (1.0 + 0.01 * (i % 11))).contiguous().to(DEV), torch.tensor([T0, T0 - 300]).to(DEV))

    (a, sa), (b, sb) = fresh(), fresh()
    plan = a[0].splan
    spans = sorted((plan.offsets[id(p)], p.numel(), n) for n, p in a[2].named_parameters() if id(p) in plan.offsets)

    def name_of(off):
        for o, n, nm in spans:
            if o <= off < o + n:
                return "%s[%d]" % (nm, off - o)
        return "arena[%d] (gap / tail; total %d)" % (off, plan.total)

    for i, T0 in enumerate(shapes):
        wave, ln = batch(i, T0)
        ops.rng_set(100 + i)
        a[0].run_step(wave, ln)
        ops.rng_set(100 + i)
        ncached = len(b[0]._gcache)
        b[0].step(wave, ln)
        torch.cuda.synchronize()
        la, lb = a[0].loss_terms()["loss"], b[0].loss_terms()["loss"]
        ga, gb = float(a[0].gnorm.sqrt()), float(b[0].gnorm.sqrt())
        na = float(torch.sqrt(sum((t.double() ** 2).sum() for t in sa)))
        nb = float(torch.sqrt(sum((t.double() ** 2).sum() for t in sb)))
        d0 = (sa[0] - sb[0]).abs()
        d1 = (sa[1] - sb[1]).abs()
        gmax = float(sa[0].abs().max())
        k = int(d0.argmax())
        flag = " <<<" if abs(ga - gb) > 1e-5 * ga else ""
        print("step %2d T0=%d graphs=%d->%d dloss=%.1e gnorm %.6f/%.6f (d %.1e) snapshot-norm %.6f/%.6f student-max-diff %.2e "
              "at %s heads-max-diff %.2e%s" % (i, T0, ncached, len(b[0]._gcache), abs(la - lb) / abs(la), ga, gb,
                                               abs(ga - gb) / ga, na, nb, float(d0.max()) / gmax, name_of(k),
                                               float(d1.max()) / gmax, flag))
        if flag:
            top = torch.topk(d0, 6)
            for v, idx in zip(top.values.tolist(), top.indices.tolist()):
                print("        %.3e  %s  eager %.5e graph %.5e" % (v / gmax, name_of(idx), float(sa[0][idx]), float(sb[0][idx])))


def first_replay(args, g):
    ops.manual_seed(21)
    ops.rng_set(0)
    s = build_step(g, DEV, torch.float32, learning_rate=0.0, reg_learning_rate=0.0)
    for h in s[2].modules():
        if hasattr(h, "noise"):
            h.noise = None
    st = s[0]
    if args.no_streams:
        st.two_streams = False
    gsnap = [torch.zeros_like(t) for t in st._grad_buffers()]
    orig = st._optimizer_step

    def spy():
        for d, t in zip(gsnap, st._grad_buffers()):
            d.copy_(t)
        orig()

    st._optimizer_step = spy
    shapes = [8000, 5000, 8000, 5000, 6500, 8000, 5000, 6500, 8000, 6500, 5000]
    # copies of the loss backward's inputs, taken (stream-ordered / inside the graph) right before it runs
    dbg = {}
    orig_bwd = ops.distill_loss_bwd

    def bwd_spy(pred, tgt, rowstat, up, dpred, *rest):
        k = len([n for n in dbg if n.startswith("<in:pred")])
        for nm, t in (("pred", pred), ("tgt", tgt), ("rowstat", rowstat)):
            key = "<in:%s#%d>" % (nm, k)
            dbg[key] = torch.empty(t.numel() + 4096, dtype=t.dtype, device=t.device)[: t.numel()].view(t.shape)
            dbg[key].copy_(t)
        return orig_bwd(pred, tgt, rowstat, up, dpred, *rest)

    if args.spy:
        ops.distill_loss_bwd = bwd_spy

    def batch(i, T0):
        return ((g["wave"][:, :T0] * (1.0 + 0.01 * (i % 11))).contiguous().to(DEV), torch.tensor([T0, T0 - 300]).to(DEV))

    def snapshot():
        torch.cuda.synchronize()
        out = {}
        for name, v in dbg.items():
            out[name] = v.detach().clone()
        for al in (st.At, st.As, st.Ah):
            for name, v in al.views.items():
                out[name] = v.detach().clone()
        out["<grads:student>"] = gsnap[0].clone()
        out["<grads:heads>"] = gsnap[1].clone()
        return out

    for i, T0 in enumerate(shapes):
        wave, ln = batch(i, T0)
        ops.rng_set(100 + i)
        n0 = len(st._gcache)
        dbg.clear()
        st.step(wave, ln)
        if len(st._gcache) == n0:
            continue
        first = snapshot()
        key = (tuple(wave.shape), True)
        gr, w_in, l_in = st._gcache[key]
        reps = []
        for _ in range(2):
            ops.rng_set(100 + i)
            w_in.copy_(wave)
            l_in.copy_(ln)
            st._push_scalars()
            gr.replay()
            reps.append(snapshot())
        print("step %d T0=%d: graph captured; buffers differing between replay #1 and #2 (and #2 vs #3):" % (i, T0))
        nd = 0
        for name in first:
            a, b, c = first[name].double(), reps[0][name].double(), reps[1][name].double()
            if a.shape != b.shape:
                continue
            d12 = float((a - b).abs().max()) if a.numel() else 0.0
            d23 = float((b - c).abs().max()) if a.numel() else 0.0
            sc = max(float(b.abs().max()), 1e-30) if a.numel() else 1.0
            nan = bool(torch.isnan(a).any()) or bool(torch.isnan(b).any())
            if d12 > 1e-6 * sc or d23 > 1e-6 * sc or nan:
                nd += 1
                print("    %-28s shape %-22s rel diff #1/#2 %.2e   #2/#3 %.2e%s" % (name, tuple(a.shape), d12 / sc, d23 / sc,
                                                                                  "  NaN" if nan else ""))
                if name.startswith("<in:") or name.startswith("h_d") or name in ("s_hc_dmasks", "s_L0_ln2_dx"):
                    d = (a - b).abs().reshape(a.shape[0], -1).max(1).values if a.dim() > 1 else (a - b).abs()
                    rows = [int(r) for r in torch.nonzero(d > 1e-6 * sc).flatten()[:24]]
                    print("        rows/entries that differ: %s" % rows)
        print("    (%d of %d buffers)" % (nd, len(first)))


if __name__ == "__main__":
    main()
