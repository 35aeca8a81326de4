"""CUDA kernels of the CTR operator family (csrc/cuda/ctr_ops.cu + ops/tc_linear.py) against the
plain PyTorch fp32 statements of the same ops (``ops/ctr_ops.py::*_torch`` / the CPU paths)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from paddlebox_b200.ops import ctr_ops as CO


def dev():
    return torch.device("cuda", 0)


def _r16(t):
    return t.to(torch.bfloat16).float()


def test_rank_attention_kernel_fwd_bwd():
    g = torch.Generator().manual_seed(3)
    n, c, R, out = 37, 24, 3, 40
    sid = torch.randint(0, 12, (n,), generator=g).sort().values
    rk = torch.randint(0, 5, (n,), generator=g)                       # ranks 0 and 4 are invalid
    ro = CO.build_rank_offset(sid, rk, R)
    x = torch.randn(n, c, generator=g)
    P = torch.randn(R * R * c, out, generator=g) * 0.2
    go = torch.randn(n, out, generator=g)
    # oracle on bf16-rounded operands (the kernel's operand type), fp32 math
    xo, Po = _r16(x).requires_grad_(True), _r16(P).requires_grad_(True)
    yo = CO.rank_attention_torch(xo, ro, Po, R)
    yo.backward(go)
    xg, Pg = x.to(dev()).requires_grad_(True), P.to(dev()).requires_grad_(True)
    y = CO.rank_attention(xg, ro.to(dev()), Pg, R)
    assert y.is_cuda and torch.allclose(y.cpu(), yo.detach(), rtol=2e-3, atol=2e-3)
    y.backward(go.to(dev()))
    # gradients: dOut is rounded to bf16 by the kernels' operand cast
    assert torch.allclose(Pg.grad.cpu(), Po.grad, rtol=2e-2, atol=3e-2)
    assert torch.allclose(xg.grad.cpu(), xo.grad, rtol=2e-2, atol=3e-2)


@pytest.mark.parametrize("layout", ["default", "batchcount", "transpose"])
def test_batch_fc_on_grouped_gemm(layout):
    g = torch.Generator().manual_seed(5)
    S, ins, fin, fout = 5, 33, 20, 12
    if layout == "default":
        x = torch.randn(S, ins, fin, generator=g)
        w = torch.randn(S, fin, fout, generator=g) * 0.3
        b = torch.randn(S, fout, generator=g)
        kw = {}
    elif layout == "batchcount":
        x = torch.randn(ins, S * fin, generator=g)
        w = torch.randn(fin, S * fout, generator=g) * 0.3
        b = torch.randn(1, S * fout, generator=g)
        kw = dict(batchcount=S)
    else:
        x = torch.randn(S, ins, fin, generator=g)
        w = torch.randn(fin, S * fout, generator=g) * 0.3
        b = torch.randn(1, S * fout, generator=g)
        kw = dict(batchcount=S, transpose_weight=True)
    xo, wo, bo = _r16(x).requires_grad_(True), _r16(w).requires_grad_(True), b.clone().requires_grad_(True)
    yo = CO.batch_fc(xo, wo, bo, "relu", **kw)
    go = torch.randn(yo.shape, generator=g)
    yo.backward(go)
    xg, wg, bg = (t.to(dev()).requires_grad_(True) for t in (x, w, b))
    y = CO.batch_fc(xg, wg, bg, "relu", **kw)
    assert y.shape == yo.shape and torch.allclose(y.cpu(), yo.detach(), rtol=2e-3, atol=2e-3)
    y.backward(go.to(dev()))
    assert torch.allclose(xg.grad.cpu(), xo.grad, rtol=2e-2, atol=3e-2)
    assert torch.allclose(wg.grad.cpu(), wo.grad, rtol=2e-2, atol=6e-2)
    assert torch.allclose(bg.grad.cpu(), bo.grad, rtol=1e-3, atol=1e-3)


def test_scaled_fc_and_int8fc_kernels():
    g = torch.Generator().manual_seed(7)
    M, K, Nn = 50, 36, 28
    x, w, b = torch.randn(M, K, generator=g), torch.randn(K, Nn, generator=g) * 0.2, torch.randn(Nn, generator=g)
    go = torch.randn(M, Nn, generator=g)
    for fn, kw, tol in [(CO.scaled_fc, dict(input_scale_factor=4.0, bias_scale_factor=2.0, grad_scale_factor=64.0), 3e-2),
                        (CO.scaled_int8fc, dict(input_expand_factor=3.0, weight_expand_factor=2.0), 3e-2),
                        (CO.scaled_int8fc, dict(dequant_weights=True), 3e-2)]:
        xo, wo, bo = (t.clone().requires_grad_(True) for t in (x, w, b))
        yo = fn(xo, wo, bo, **kw)
        yo.backward(go)
        xg, wg, bg = (t.to(dev()).requires_grad_(True) for t in (x, w, b))
        y = fn(xg, wg, bg, **kw)
        scale = float(yo.detach().abs().max()) + 1.0
        assert torch.allclose(y.cpu(), yo.detach(), rtol=3e-3, atol=3e-3 * scale), fn.__name__
        y.backward(go.to(dev()))
        for a, o in ((xg, xo), (wg, wo), (bg, bo)):
            assert torch.allclose(a.grad.cpu(), o.grad, rtol=tol, atol=tol * (float(o.grad.abs().max()) + 1)), fn.__name__


def _bags(B, S, max_len, gen):
    lens = torch.randint(0, max_len + 1, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    return lod


@pytest.mark.parametrize("cfg", [dict(), dict(use_cvm=False), dict(need_filter=True, threshold=0.6),
                                 dict(quant_ratio=64, pad_value=0.25), dict(cvm_offset=6, max_cvm_offset=8)])
def test_seqpool_pcoc_kernel(cfg):
    g = torch.Generator().manual_seed(11)
    B, S, emb = 19, 3, 5
    used = cfg.get("cvm_offset", 7)
    mx = cfg.get("max_cvm_offset", used)
    H = mx + emb
    lod = _bags(B, S, 4, g)
    n = int(lod[-1])
    x = torch.rand(n + 2, H, generator=g) * 2                      # two trailing rows beyond lod[-1]
    cvm = torch.rand(B, 4, generator=g)
    q = torch.rand(B, used - 4, generator=g)
    xo = x.clone().requires_grad_(True)
    yo = CO.fused_seqpool_cvm_with_pcoc(xo, lod, B, S, cvm=cvm, q_values=q, **cfg)
    go = torch.randn(yo.shape, generator=g)
    yo.backward(go)
    xg = x.to(dev()).requires_grad_(True)
    y = CO.fused_seqpool_cvm_with_pcoc(xg, lod.to(dev()), B, S, cvm=cvm.to(dev()), q_values=q.to(dev()), **cfg)
    assert y.is_cuda and torch.allclose(y.cpu(), yo.detach(), rtol=1e-5, atol=1e-5)
    y.backward(go.to(dev()))
    assert torch.allclose(xg.grad.cpu()[:n], xo.grad[:n], rtol=1e-6, atol=1e-6)
    assert float(xg.grad[n:].abs().sum()) == 0.0


def test_fused_concat_and_seqpool_concat_kernels():
    g = torch.Generator().manual_seed(13)
    B = 23
    xs = [torch.randn(B, w, generator=g) for w in (9, 12, 9, 15)]
    xo = [t.clone().requires_grad_(True) for t in xs]
    yo = CO.fused_concat(xo, start_index=2, length=6)
    go = torch.randn(yo.shape, generator=g)
    yo.backward(go)
    xg = [t.to(dev()).requires_grad_(True) for t in xs]
    y = CO.fused_concat(xg, start_index=2, length=6)
    assert y.is_cuda and torch.equal(y.cpu(), yo.detach())
    y.backward(go.to(dev()))
    for a, o in zip(xg, xo):
        assert torch.equal(a.grad.cpu(), o.grad)
    # pooled variant: two groups sharing one lod
    S, Bp = 4, 17
    lod = _bags(Bp, S, 3, g)
    n = int(lod[-1])
    x1, x2 = torch.randn(n + 1, 7, generator=g), torch.randn(n + 1, 5, generator=g)
    o1, o2 = x1.clone().requires_grad_(True), x2.clone().requires_grad_(True)
    yo = CO.fused_seqpool_concat([o1, o2], S, offsets=[1, 0], dims=[4, 0], lod=lod, batch_size=Bp)
    go = torch.randn(yo.shape, generator=g)
    yo.backward(go)
    g1, g2 = x1.to(dev()).requires_grad_(True), x2.to(dev()).requires_grad_(True)
    y = CO.fused_seqpool_concat([g1, g2], S, offsets=[1, 0], dims=[4, 0], lod=lod.to(dev()), batch_size=Bp)
    assert y.shape == yo.shape and torch.allclose(y.cpu(), yo.detach(), rtol=1e-5, atol=1e-5)
    y.backward(go.to(dev()))
    assert torch.allclose(g1.grad.cpu(), o1.grad, atol=1e-6) and torch.allclose(g2.grad.cpu(), o2.grad, atol=1e-6)


@pytest.mark.parametrize("ad_off", [0, 1])
def test_fused_seq_tensor_kernel(ad_off):
    g = torch.Generator().manual_seed(17)
    ins, bc, L, sn, d, an = 5, 3, 6, 4, 3, 2
    xs = torch.rand(ins, bc * sn * L * d, generator=g) + 0.1
    xs.view(ins, bc, sn, L, d)[1, 2, :, 3:] = 0
    ad = torch.rand(ins, bc * an * d, generator=g)
    want = CO.fused_seq_tensor(xs, ad, bc, L, sn, d, an, ad_off)
    got = CO.fused_seq_tensor(xs.to(dev()), ad.to(dev()), bc, L, sn, d, an, ad_off)
    for a, o, nm in zip(got, want, ("din", "mask", "side", "sess")):
        assert a.is_cuda and a.shape == o.shape and torch.allclose(a.cpu(), o, atol=1e-6), nm


def test_cross_norm_hadamard_kernel():
    g = torch.Generator().manual_seed(19)
    B, f, d = 29, 3, 4
    x = torch.randn(B, f * 2 * d, generator=g)
    torch.manual_seed(0)
    mo = CO.CrossNormHadamard(f, d)
    mg = CO.CrossNormHadamard(f, d).to(dev())
    mg.load_state_dict(mo.state_dict())
    xo, xg = x.clone().requires_grad_(True), x.to(dev()).requires_grad_(True)
    yo, y = mo(xo), mg(xg)
    assert y.is_cuda and torch.allclose(y.cpu(), yo.detach(), rtol=1e-4, atol=1e-5)
    go = torch.randn(yo.shape, generator=g)
    yo.backward(go)
    y.backward(go.to(dev()))
    assert torch.allclose(xg.grad.cpu(), xo.grad, rtol=1e-4, atol=1e-5)
    for nm in ("batch_size", "batch_sum", "batch_square_sum"):          # the summaries moved the same way
        assert torch.allclose(getattr(mg.norm, nm).cpu(), getattr(mo.norm, nm), rtol=1e-4, atol=1e-4), nm


@pytest.mark.parametrize("nranks", [2, 4])
def test_data_norm_stat_update_virtual_ranks(nranks):
    """3 statistics vectors summed over the ranks + decayed summary update in one launch."""
    from paddlebox_b200.parallel.symm import SymmetricBuffer
    g = torch.Generator().manual_seed(23)
    Cn, decay = 37, 0.75
    bufs = SymmetricBuffer.emulated_group(nranks, 3 * Cn + 8)
    stats = [torch.rand(3, Cn, generator=g) for _ in range(nranks)]
    summ = [[torch.rand(Cn, generator=g).to(dev()) for _ in range(3)] for _ in range(nranks)]
    before = [[t.clone().cpu() for t in trio] for trio in summ]
    for b, s in zip(bufs, stats):
        b.tensor[:3 * Cn].copy_(s.reshape(-1).to(dev()))
    SymmetricBuffer.stat_update_emu(bufs, 0, Cn, 1.0 / nranks, decay, summ)
    mean = torch.stack(stats).sum(0) / nranks
    for r in range(nranks):
        for j in range(3):
            assert torch.allclose(summ[r][j].cpu(), before[r][j] * decay + mean[j], rtol=1e-6, atol=1e-6), (r, j)
