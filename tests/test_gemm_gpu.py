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


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("b_mn", [0, 1])
@pytest.mark.parametrize("M,N,K", [(3992, 3072, 768), (500, 320, 192)])
def test_fused_activation_backward(cg, b_mn, M, N, K):
    """dgrad GEMM whose epilogue differentiates  o = colmask * GELU(z)  (EPI_BWD): dz, bias and mask column sums against
    the restatement; ragged M (rows beyond M contribute nothing to the column sums)."""
    a, b = _views(M, N, K, 0, b_mn)
    z = rnd(M, N, dtype=BF16)
    colmask = torch.rand(N, device=DEV) * 0.8 + 0.2
    c = zer(M, N, dtype=BF16)
    dbias, dmask = rnd(N), rnd(N)   # accumulated into
    both("gemm", (a, b, c, M, N, K), dict(act=1, colmask=colmask, act_bwd=True, z=z, dbias=dbias, dmask=dmask,
                                          cta_group=cg), fn=G.gemm, ref=emu.gemm, tol=2e-2)


def test_fused_activation_backward_strided_batched():
    """the conv-stack use: per-utterance batches, output rows interleaved by phase into a zero-padded row layout,
    pre-activation read with its own strides"""
    B, rows, C, K, s, ph, padr, Tin = 3, 130, 128, 256, 2, 1, 3, 261
    rows_pb = Tin + 2 * padr
    a = rnd(B, rows, K, dtype=BF16)
    w = rnd(C, K, dtype=BF16, scale=0.05)
    z = rnd(B, Tin, C, dtype=BF16)
    out = zer(B, rows_pb, C, dtype=BF16)
    colmask = torch.rand(C, device=DEV)
    dbias, dmask = zer(C), zer(C)
    av = G.View(a, rows, K, K, sb0=rows * K)
    both("gemm", (av, G.mat(w), out, rows, C, K),
         dict(nb0=B, ldc=s * C, c_sb0=rows_pb * C, c_offset=(padr + ph) * C, act=1, colmask=colmask, act_bwd=True, z=z,
              ldz=s * C, z_sb0=Tin * C, z_offset=ph * C, dbias=dbias, dmask=dmask), fn=G.gemm, ref=emu.gemm, tol=2e-2)


def test_fused_activation_backward_dropout_mask_matches_forward():
    """with dropout: the backward epilogue regenerates the forward GEMM's keep mask (same seed, same element index)"""
    M, F, D, p, seed = 1000, 3072, 768, 0.1, 777
    x, w1, bias = rnd(M, D, dtype=BF16), rnd(F, D, dtype=BF16, scale=0.05), rnd(F)
    h0, h1, z = zer(M, F, dtype=BF16), zer(M, F, dtype=BF16), zer(M, F, dtype=BF16)
    G.gemm(G.mat(x), G.mat(w1), h0, M, F, D, bias=bias, act=1, c2=z)
    G.gemm(G.mat(x), G.mat(w1), h1, M, F, D, bias=bias, act=1, drop_p=p, seed=seed, c2=z)
    g, w2 = rnd(M, D, dtype=BF16), rnd(D, F, dtype=BF16, scale=0.05)
    dz0, dz1 = zer(M, F, dtype=BF16), zer(M, F, dtype=BF16)
    G.gemm(G.mat(g), G.mat_t(w2), dz0, M, F, D, act=1, act_bwd=True, z=z)
    G.gemm(G.mat(g), G.mat_t(w2), dz1, M, F, D, act=1, drop_p=p, seed=seed, act_bwd=True, z=z)
    torch.cuda.synchronize()
    big = (h0.float().abs() > 0.05 * h0.float().abs().mean()) & (dz0.float().abs() > 0.05 * dz0.float().abs().mean())
    fwd_keep = h1.float().abs() > 0
    bwd_keep = dz1.float().abs() > 0
    assert torch.equal(fwd_keep[big], bwd_keep[big])
    ratio = (dz1.float() / dz0.float())[big & bwd_keep]
    assert ((ratio * (1 - p) - 1).abs() < 2e-2).all()


@pytest.mark.parametrize("bn", [0, 256, 128])
def test_taps_folded_into_n(bn):
    """grouped-conv weight gradient with the taps folded into N (b_nsplit): column n = tap * 64 + channel reads the
    activation view shifted by `tap` rows; the restatement runs one product per tap"""
    Gn, Cgp, K, Text = 3, 64, 16, 700
    Wd = Gn * Cgp
    front = K // 2
    rows = front + Text + K
    dzp = rnd(rows, Wd, dtype=BF16)
    xp = rnd(rows, Wd, dtype=BF16)
    dwp = zer(Gn, Cgp, K, Cgp)
    a = G.View(dzp, Cgp, Text, 1, cs=Wd, sb0=Cgp, offset=front * Wd)
    bv = G.View(xp, Cgp, Text, 1, cs=Wd, sb0=Cgp, sb1=Wd)
    both("gemm", (a, bv, dwp, Cgp, K * Cgp, Text), dict(nb0=Gn, ldc=K * Cgp, c_sb0=Cgp * K * Cgp, b_taps=K, cta_group=1,
                                                       block_n=bn), fn=G.gemm, ref=emu.gemm, tol=2e-3)
    # and it equals the per-tap batched formulation it replaces
    ref = zer(Gn, Cgp, K, Cgp)
    G.gemm(a, bv, ref, Cgp, Cgp, Text, nb0=Gn, nb1=K, ldc=K * Cgp, c_sb0=Cgp * K * Cgp, c_sb1=Cgp)
    torch.cuda.synchronize()
    assert (dwp - ref).abs().max().item() < 1e-3 * ref.abs().max().item()
