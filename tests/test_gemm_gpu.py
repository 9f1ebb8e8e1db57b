"""The tcgen05 GEMM (`dph_gemm_bf16`) against the torch restatement in tests/emu.py: operand majors, tile widths, one CTA
per tile and CTA pairs (cta_group::2), fused epilogues, split-K atomics, strided-conv and batched views."""
import pytest
import torch

import emu
from dphubert_b200 import gemm as G
from test_kernels_gpu import both, rnd, zer, F32, BF16, DEV  # noqa: F401

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _seed():
    torch.manual_seed(4321)


def _views(M, N, K, a_mn, b_mn):
    a = rnd(K, M, dtype=BF16) if a_mn else rnd(M, K, dtype=BF16)
    b = rnd(K, N, dtype=BF16) if b_mn else rnd(N, K, dtype=BF16)
    return (G.mat_t(a) if a_mn else G.mat(a)), (G.mat_t(b) if b_mn else G.mat(b))


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("bn", [64, 128, 192, 256])
@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_majors_and_tiles(cg, bn, a_mn, b_mn):
    if cg == 2 and b_mn and bn % 128:
        pytest.skip("an MN-major B tile of 64/192 columns cannot be halved into whole TMA boxes")
    M, N, K = 520, 392, 328  # ragged in every dimension (multiples of 8 keep the TMA strides legal)
    a, b = _views(M, N, K, a_mn, b_mn)
    c = zer(M, N)
    both("gemm", (a, b, c, M, N, K), dict(block_n=bn, cta_group=cg), fn=G.gemm, ref=emu.gemm, tol=2e-3)


@pytest.mark.parametrize("cg", [0, 1, 2])
@pytest.mark.parametrize("M,N,K", [(3992, 768, 768), (100, 72, 40), (129, 520, 1000), (1000, 2304, 768), (257, 64, 64)])
def test_sizes_auto_tile(cg, M, N, K):
    a, b = _views(M, N, K, 0, 0)
    c = zer(M, N, dtype=BF16)
    both("gemm", (a, b, c, M, N, K), dict(cta_group=cg), fn=G.gemm, ref=emu.gemm, tol=2e-2)


@pytest.mark.parametrize("cg", [1, 2])
def test_fused_epilogue(cg):
    M, N, K = 500, 320, 192
    a, b = _views(M, N, K, 0, 0)
    bias, colmask, smask = rnd(N), (torch.rand(N, device=DEV) > 0.3).float() * 1.7, rnd(1)
    res = rnd(M, N, dtype=BF16)
    c, c2 = zer(M, N, dtype=BF16), zer(M, N, dtype=BF16)
    both("gemm", (a, b, c, M, N, K), dict(bias=bias, act=1, colmask=colmask, scalar_mask=smask, residual=res, c2=c2,
                                          alpha=0.5, cta_group=cg), fn=G.gemm, ref=emu.gemm, tol=2e-2)


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("ksplit", [1, 7, -1])
def test_wgrad_split_k_atomic(cg, ksplit):
    M, N, K = 768, 520, 3992
    a, b = _views(M, N, K, 1, 1)
    c = rnd(M, N)  # accumulated into
    both("gemm", (a, b, c, M, N, K), dict(atomic=True, ksplit=ksplit, cta_group=cg), fn=G.gemm, ref=emu.gemm, tol=2e-3)


@pytest.mark.parametrize("cg", [1, 2])
def test_aliased_batches_accumulate(cg):
    """conv weight gradient: every utterance (batch index) adds into the same output."""
    B, M, N, K = 3, 256, 384, 200
    a = rnd(B, K, M, dtype=BF16)
    b = rnd(B, K, N, dtype=BF16)
    av = G.View(a, M, K, 1, cs=M, sb0=K * M)
    bv = G.View(b, N, K, 1, cs=N, sb0=K * N)
    c = zer(M, N)
    both("gemm", (av, bv, c, M, N, K), dict(nb0=B, c_sb0=0, atomic=True, ksplit=-1, cta_group=cg), fn=G.gemm,
         ref=emu.gemm, tol=2e-3)


@pytest.mark.parametrize("cg", [1, 2])
def test_strided_conv_view(cg):
    """Conv1d(C, C, k=3, stride=2) as an implicit GEMM over overlapping rows of the channels-last activation."""
    B, T, C, k, s = 2, 301, 128, 3, 2
    To = (T - k) // s + 1
    x = rnd(B, T, C, dtype=BF16)
    w = rnd(C, k * C, dtype=BF16, scale=0.05)
    av = G.View(x, To, k * C, s * C, sb0=T * C)
    y = zer(B, To, C, dtype=BF16)
    both("gemm", (av, G.mat(w), y, To, C, k * C), dict(nb0=B, c_sb0=To * C, act=1, cta_group=cg), fn=G.gemm,
         ref=emu.gemm, tol=2e-2)
