"""Correctness and speed of `dph_multimem_allreduce` (csrc/allreduce.cu) against NCCL on N GPUs of one box:

    torchrun --nproc-per-node N --master-addr 127.0.0.1 profiles/multimem_check.py

Every rank fills a symmetric-memory buffer with a rank-dependent pattern, reduces slices of it over the NVSwitch
multicast address and compares with the closed form; then times both paths on the step's slice sizes (CUDA events,
max over ranks)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dphubert_b200 import ops  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    import torch.distributed._symmetric_memory as symm_mem

    n = 96 * 1024 * 1024   # 384 MB of fp32: the student's gradient arena
    buf = symm_mem.empty(n, dtype=torch.float32, device=dev)
    hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
    if rank == 0:
        print("multicast_ptr %#x, world %d" % (hdl.multicast_ptr, hdl.world_size), flush=True)
    assert hdl.multicast_ptr, "no multicast support"
    idx = torch.arange(n, device=dev, dtype=torch.float32)

    def fill():
        buf.copy_((idx % 1021.0) * 1e-3 + float(rank + 1))

    def mm(off, cnt, ctas=16):
        hdl.barrier(0)
        ops.multimem_allreduce(hdl.multicast_ptr + 4 * off, cnt, rank, world, 1.0 / world, ctas)
        hdl.barrier(0)

    # ---- correctness: two slices, the rest must stay untouched
    fill()
    torch.cuda.synchronize()
    mm(0, 7 * 1024 * 1024)
    mm(50 * 1024 * 1024 + 4, 1000 * 4)
    torch.cuda.synchronize()
    want = (idx % 1021.0) * 1e-3 + (world + 1) / 2.0
    own = (idx % 1021.0) * 1e-3 + float(rank + 1)
    ok = True
    for lo, hi, ref in ((0, 7 * 1024 * 1024, want), (7 * 1024 * 1024, 50 * 1024 * 1024 + 4, own),
                        (50 * 1024 * 1024 + 4, 50 * 1024 * 1024 + 4004, want), (50 * 1024 * 1024 + 4004, n, own)):
        err = float((buf[lo:hi] - ref[lo:hi]).abs().max())
        ok = ok and err < 1e-5
        if rank == 0:
            print("slice [%d, %d): max err %.2e" % (lo, hi, err), flush=True)
    flag = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    assert float(flag) == 1.0, "multimem all-reduce mismatch"

    # ---- speed: one transformer layer's slice (28 MB), three layers (85 MB), the whole arena (384 MB)
    def timeit(fn, reps=10):
        for _ in range(3):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    plain = torch.zeros(n, device=dev)
    for cnt in (7 * 1024 * 1024, 21 * 1024 * 1024, n):
        row = []
        for ctas in (4, 8, 16, 32):
            row.append(timeit(lambda: mm(0, cnt, ctas)))
        t_nccl = timeit(lambda: dist.all_reduce(plain[:cnt], op=dist.ReduceOp.AVG))
        if rank == 0:
            gb = cnt * 4 / 1e9
            print("%6.1f MB: multimem %s ms (4 / 8 / 16 / 32 CTAs) = %s GB/s algorithmic;  NCCL %.3f ms = %.0f GB/s" % (
                cnt * 4 / 1e6, " / ".join("%.3f" % x for x in row), " / ".join("%.0f" % (gb / x * 1e3) for x in row), t_nccl,
                gb / t_nccl * 1e3), flush=True)
    # ---- inside a CUDA graph
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    torch.cuda.synchronize()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            mm(0, 7 * 1024 * 1024)
    fill()
    torch.cuda.synchronize()
    g.replay()
    torch.cuda.synchronize()
    err = float((buf[: 7 * 1024 * 1024] - want[: 7 * 1024 * 1024]).abs().max())
    if rank == 0:
        print("captured graph replay: max err %.2e" % err, flush=True)
    assert err < 1e-5
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
