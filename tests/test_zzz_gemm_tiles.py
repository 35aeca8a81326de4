"""tcgen05 GEMM with the tile width FORCED (auto selection only ever picked 64, and 128 on split-K fp32, in the
other tests): every epilogue — fp32, bias + ReLU with bf16 row-major and transposed stores, ReLU-backward mask
with alpha, split-K accumulate, bias-gradient column — at BN = 32 / 64 / 128 / 256, on shapes with partial
tiles in both directions, against an fp32 PyTorch product of the same bf16 operands."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda", 0)


@pytest.mark.parametrize("block_n", [32, 64, 128, 256])
@pytest.mark.parametrize("M,N,K", [(300, 200, 136), (512, 520, 264), (4096, 512, 496)])
def test_gemm_forced_tile_width_all_epilogues(block_n, M, N, K):
    from paddlebox_b200.ops.mlp import _pad8, gemm_bf16_tn
    torch.manual_seed(M + N + K + block_n)
    A = torch.zeros(M, _pad8(K), dtype=torch.bfloat16, device=dev())
    Bm = torch.zeros(N, _pad8(K), dtype=torch.bfloat16, device=dev())
    A[:, :K] = torch.randn(M, K, device=dev()) * 0.5
    Bm[:, :K] = torch.randn(N, K, device=dev()) * 0.5
    bias = torch.randn(N, device=dev())
    ref = A[:, :K].float() @ Bm[:, :K].float().t()
    tol = dict(rtol=2e-2, atol=2e-2 * math.sqrt(K / 64))
    kw = dict(impl=0, block_n=block_n)
    out = torch.full((M, N), float("nan"), device=dev())
    gemm_bf16_tn(A, Bm, M, N, K, out_f32=out, **kw)
    assert torch.allclose(out, ref, **tol), float((out - ref).abs().max())
    # bias + ReLU, bf16 row-major + transposed stores
    ob = torch.zeros(M, _pad8(N), dtype=torch.bfloat16, device=dev())
    ot = torch.zeros(N, _pad8(M), dtype=torch.bfloat16, device=dev())
    gemm_bf16_tn(A, Bm, M, N, K, bias=bias, out_bf16=ob, out_t=ot, relu=True, **kw)
    r2 = torch.relu(ref + bias)
    assert torch.allclose(ob[:, :N].float(), r2, rtol=3e-2, atol=tol["atol"] + 2e-2)
    assert torch.equal(ot[:, :M], ob[:, :N].t())
    # ReLU-backward mask and alpha
    mask = torch.zeros(M, _pad8(N), dtype=torch.bfloat16, device=dev())
    mask[:, :N] = torch.randn(M, N, device=dev())
    om = torch.full((M, N), float("nan"), device=dev())
    gemm_bf16_tn(A, Bm, M, N, K, mask=mask, out_f32=om, alpha=0.5, **kw)
    assert torch.allclose(om, 0.5 * ref * (mask[:, :N].float() > 0), **tol)
    # split-K accumulate
    acc = torch.ones(M, N, device=dev())
    gemm_bf16_tn(A, Bm, M, N, K, out_f32=acc, accumulate=True, splitk=3, **kw)
    assert torch.allclose(acc, ref + 1, **tol)
    # bias-gradient column routed to a separate vector
    acc2 = torch.zeros(M, N - 1, device=dev())
    colv = torch.zeros(M, device=dev())
    gemm_bf16_tn(A, Bm, M, N, K, out_f32=acc2, accumulate=True, splitk=2, col_out=colv, col_split=N - 1, **kw)
    assert torch.allclose(acc2, ref[:, :N - 1], **tol) and torch.allclose(colv, ref[:, N - 1], **tol)
