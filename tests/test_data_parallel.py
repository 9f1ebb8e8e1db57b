"""The N > 1 path on CPU: two `gloo` ranks, each running the native step (kernel library emulated, tests/emu.py) on its
own utterance of the golden batch.  The per-layer overlapped gradient all-reduce must reproduce the reference's
full-batch gradients (DDP semantics, distill.py:41), and the ranks must stay bit-identical after the optimizer step."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, case, q):
    try:
        for p in (os.path.dirname(HERE), os.path.join(os.path.dirname(HERE), "oracle"), HERE):
            if p not in sys.path:
                sys.path.insert(0, p)
        torch.set_num_threads(2)
        import conftest
        import emu
        import helpers

        mpatch = pytest.MonkeyPatch()
        emu.install(mpatch)
        dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
        g = conftest.load_golden(case)
        dev = torch.device("cpu")
        step, teacher, student, projs, lam = helpers.build_step(g, dev, torch.float32, process_group=dist.group.WORLD)
        assert step.world == world and step.splan.grad_hook is not None
        wave, lengths = g["wave"][rank:rank + 1].contiguous(), g["lengths"][rank:rank + 1].contiguous()
        step._push_scalars()
        step._body(wave, lengths)
        n_layer_reduces = len(step._pending)
        step._allreduce()
        errs = helpers.grad_errors(g, student, projs, lam)
        worst = max(errs.values())
        # full step: parameters must stay identical on every rank
        step._optimizer_step()
        digest = torch.stack([a.double().sum() for a, _ in step.arenas()] + [a.double().abs().sum() for a, _ in step.arenas()])
        gathered = [torch.zeros_like(digest) for _ in range(world)]
        dist.all_gather(gathered, digest)
        same = all(torch.equal(gathered[0], x) for x in gathered)
        q.put((rank, worst, n_layer_reduces, same, None))
        dist.destroy_process_group()
        mpatch.undo()
    except Exception as e:  # noqa: BLE001
        import traceback

        q.put((rank, None, None, None, traceback.format_exc() + repr(e)))


@pytest.mark.parametrize("ar_group", ["1", "2"])
@pytest.mark.parametrize("case", ["step_base"])
def test_two_rank_gradient_mean_matches_full_batch_reference(case, ar_group, monkeypatch):
    # transformer layers per collective (trainer.ar_group): one slice per layer / both layers of the fixture in one
    monkeypatch.setenv("DPH_AR_GROUP", ar_group)
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, worst, n_layer, same, err in results:
        assert err is None, "rank %d failed:\n%s" % (rank, err)
        assert n_layer >= 2, "the per-layer all-reduces were not issued during the backward pass"
        assert worst < 2e-4, "rank %d: gradient mean deviates from the full-batch reference by %.3e" % (rank, worst)
        assert same, "ranks diverged after the optimizer step"
