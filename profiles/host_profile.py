"""Where does the HOST time of an eager (variable-shape) training step go?  Runs on the GPU box:
    python profiles/host_profile.py > gpurun_out/host_profile.txt
cProfile over a few `NativeDistillStep.run_step` calls of the bench workload (after warm-up), top functions by own time."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from dphubert_b200.trainer import NativeDistillStep  # noqa: E402

dev = torch.device("cuda:0")
teacher, student, projs, layers, cfg, scfg, name = bench.build_models("prune", dev, torch.bfloat16)
step = NativeDistillStep(teacher, student, layers, projs, "layer2layer", 0.0, 1.0, 1.0, "raw", use_reg=True)
wave = 0.1 * torch.randn(8, 160000, device=dev)
lengths = torch.full((8,), 160000, device=dev, dtype=torch.int64)
for _ in range(3):
    step.run_step(wave, lengths)
torch.cuda.synchronize()
N = 5
t0 = time.perf_counter()
for _ in range(N):
    step.run_step(wave, lengths)
host = (time.perf_counter() - t0) / N
torch.cuda.synchronize()
total = (time.perf_counter() - t0) / N
print("eager step: host issue %.2f ms, wall incl. GPU drain %.2f ms" % (host * 1e3, total * 1e3))
pr = cProfile.Profile()
pr.enable()
for _ in range(N):
    step.run_step(wave, lengths)
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr, stream=sys.stdout)
st.sort_stats("tottime").print_stats(28)
