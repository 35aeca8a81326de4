"""Numerics of every CUDA kernel against plain PyTorch fp32/fp64 references (B200 only)."""
import ctypes as C
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from paddlebox_b200 import _native as N
from paddlebox_b200.ops import reference as R


def dev():
    return torch.device("cuda", 0)


# ------------------------------------------------------------------ table
def _mk_table(cap=1 << 14, dim=8, opt="adagrad", **cfg):
    from paddlebox_b200.ops.sparse_table import SparseTable
    return SparseTable(cap, embedx_dim=dim, optimizer=opt, config=N.SparseOptConfig(**cfg), seed=7)


def test_table_insert_lookup_count_dump():
    t = _mk_table()
    keys = torch.randint(1, 1 << 40, (5000,), device=dev(), dtype=torch.int64).unique()
    t.insert(keys)
    t.insert(keys[:100])  # re-insert is a no-op
    assert t.size() == keys.numel()
    vals = t.lookup(keys)
    assert vals.shape == (keys.numel(), 11)
    assert torch.all(vals == 0)
    miss = t.lookup(torch.tensor([0, (1 << 50) + 12345], device=dev()))
    assert torch.all(miss == 0)
    dk, dr = t.dump()
    assert sorted(dk.tolist()) == sorted(keys.tolist())
    assert dr.shape[1] == t.layout.stride


@pytest.mark.parametrize("opt", ["adagrad", "adam", "shared_adam"])
def test_table_update_matches_oracle(opt):
    dim = 8
    t = _mk_table(dim=dim, opt=opt, mf_create_thresholds=1.0, mf_initial_range=0.0)
    L, cfg = t.layout, t.config
    keys = torch.arange(1, 65, device=dev(), dtype=torch.int64) * 7919
    t.insert(keys)
    g = torch.Generator(device="cpu").manual_seed(3)
    rows = {int(k): R.new_row(L, cfg) for k in keys.tolist()}
    for step in range(4):
        push = torch.zeros(keys.numel(), L.push_stride)
        push[:, 0] = 3
        push[:, 1] = torch.randint(1, 5, (keys.numel(),), generator=g).float()
        push[:, 2] = torch.randint(0, 2, (keys.numel(),), generator=g).float()
        push[:, 3:4 + dim] = torch.randn(keys.numel(), dim + 1, generator=g)
        t.update(keys, push.to(dev()))
        for i, k in enumerate(keys.tolist()):
            rows[k] = R.sparse_row_update(rows[k], push[i].tolist(), L, cfg)
    dk, dr = t.dump()
    got = {int(k): r for k, r in zip(dk.tolist(), dr.cpu())}
    for k in keys.tolist():
        exp = torch.tensor(rows[k])
        assert torch.allclose(got[k][:len(exp)], exp, rtol=2e-4, atol=2e-5), (k, got[k], exp)
    # embedx was created (threshold 1.0) for the keys whose score crossed it, then trained
    assert float(dr[:, L.off_mf].max()) == dim


def test_table_shrink_and_dirty():
    t = _mk_table(dim=4)
    keys = torch.arange(1, 1001, device=dev(), dtype=torch.int64)
    t.insert(keys)
    push = torch.zeros(10, t.layout.push_stride, device=dev())
    push[:, 1] = 100.0
    push[:, 2] = 50.0
    t.update(keys[:10], push)
    dk, _ = t.dump(only_dirty=True, clear_dirty=True)
    assert sorted(dk.tolist()) == list(range(1, 11))
    dk2, _ = t.dump(only_dirty=True)
    assert dk2.numel() == 0
    kept = t.shrink(decay=0.5, delete_threshold=1.0, delete_after_unseen_days=30)
    assert kept == 10 and t.size() == 10
    v = t.lookup(keys[:10])
    assert torch.allclose(v[:, 0], torch.full((10,), 50.0, device=dev()))


# ------------------------------------------------------------------ sparse pull / push
def _bags(B, S, max_len, n_keys_space, gen):
    lens = torch.randint(0, max_len + 1, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    n = int(lod[-1])
    keys = torch.randint(1, n_keys_space, (n,), generator=gen, dtype=torch.int64)
    return keys, lod


def _prime(table, keys, dim):
    """give every key non-trivial values: one push with big show so embedx is created."""
    uk = keys.unique()
    table.insert(uk)
    push = torch.zeros(uk.numel(), table.layout.push_stride, device=dev())
    push[:, 1] = 20
    push[:, 2] = 3
    push[:, 3] = torch.randn(uk.numel(), device=dev())
    table.update(uk, push)
    push[:, 3:4 + dim] = torch.randn(uk.numel(), dim + 1, device=dev())
    table.update(uk, push)
    return uk


def test_sparse_pull_push_single_rank():
    from paddlebox_b200.parallel.sparse_engine import SparseEngine
    torch.manual_seed(0)
    gen = torch.Generator().manual_seed(1)
    B, S, dim = 64, 5, 8
    H = 3 + dim
    t = _mk_table(cap=1 << 15, dim=dim)
    keys, lod = _bags(B, S, 6, 3000, gen)
    keys, lod = keys.to(dev()), lod.to(dev())
    _prime(t, keys, dim)
    eng = SparseEngine(t, max_keys=keys.numel() + 16)
    restore = eng.pull(keys)
    eng.check_status()
    got = eng.values[restore.long()][:, :H]
    exp = t.lookup(keys)
    assert torch.equal(got, exp)
    # restore indices identify unique keys
    uk, inv = torch.unique(keys, return_inverse=True)
    assert restore.unique().numel() == uk.numel()
    # ---- fused push from pooled grads vs oracle ----
    Eo = H
    dout = torch.randn(B, S * Eo, device=dev())
    cvm = torch.stack([torch.ones(B, device=dev()), (torch.rand(B, device=dev()) < 0.3).float()], 1)
    before_k, before_r = t.dump()
    eng.push_pooled(restore, dout, cvm, lod, B, S, Eo, 2, 0, -float(B))
    eng.check_status()
    after_k, after_r = t.dump()
    # oracle: per-occurrence grads -> merge -> row update
    dx = R.seqpool_cvm_bwd(dout.cpu(), cvm.cpu(), lod.cpu(), B, S, H)
    k2s = torch.repeat_interleave(torch.arange(S), (lod[1:] - lod[:-1]).cpu().view(S, B).sum(1))
    merged = R.merge_push(inv.cpu(), uk.numel(), dx, k2s, 2, -float(B))
    bmap = {int(k): r for k, r in zip(before_k.tolist(), before_r.cpu())}
    amap = {int(k): r for k, r in zip(after_k.tolist(), after_r.cpu())}
    for i, k in enumerate(uk.tolist()):
        exp_row = torch.tensor(R.sparse_row_update(bmap[k].tolist(), merged[i].tolist(), t.layout, t.config))
        # slot id of a key that occurs in several slots is "last writer wins"
        exp_row[t.layout.off_tail + 1] = amap[k][t.layout.off_tail + 1]
        assert torch.allclose(amap[k], exp_row, rtol=1e-3, atol=1e-4), (k, amap[k], exp_row)


def test_sparse_push_explicit_key_grads():
    from paddlebox_b200.parallel.sparse_engine import SparseEngine
    gen = torch.Generator().manual_seed(5)
    B, S, dim = 32, 3, 4
    H = 3 + dim
    t = _mk_table(cap=1 << 12, dim=dim)
    keys, lod = _bags(B, S, 4, 500, gen)
    keys, lod = keys.to(dev()), lod.to(dev())
    _prime(t, keys, dim)
    eng = SparseEngine(t, max_keys=keys.numel() + 8)
    restore = eng.pull(keys)
    grads = torch.randn(keys.numel(), H, device=dev())
    grads[:, 0] = 1.0
    grads[:, 1] = (torch.rand(keys.numel(), device=dev()) < 0.5).float()
    k2s = torch.empty(keys.numel(), dtype=torch.int32, device=dev())
    N.check(N.cuda().pbx_key2slot(N.ptr(lod), B, S, C.c_int64(keys.numel()), N.ptr(k2s), N.stream_ptr()))
    exp_k2s = torch.repeat_interleave(torch.arange(S), (lod[1:] - lod[:-1]).cpu().view(S, B).sum(1))
    assert torch.equal(k2s.cpu().long(), exp_k2s)
    before_k, before_r = t.dump()
    eng.push_keys(restore, grads, k2s, 2, -float(B))
    eng.check_status()
    after_k, after_r = t.dump()
    uk, inv = torch.unique(keys, return_inverse=True)
    merged = R.merge_push(inv.cpu(), uk.numel(), grads.cpu(), exp_k2s, 2, -float(B))
    bmap = {int(k): r for k, r in zip(before_k.tolist(), before_r.cpu())}
    amap = {int(k): r for k, r in zip(after_k.tolist(), after_r.cpu())}
    for i, k in enumerate(uk.tolist()):
        exp_row = torch.tensor(R.sparse_row_update(bmap[k].tolist(), merged[i].tolist(), t.layout, t.config))
        exp_row[t.layout.off_tail + 1] = amap[k][t.layout.off_tail + 1]
        assert torch.allclose(amap[k], exp_row, rtol=1e-3, atol=1e-4)


@pytest.mark.parametrize("nranks,dim,opt", [(2, 8, "adagrad"), (4, 8, "adagrad"), (8, 8, "adagrad"),
                                             (2, 64, "adagrad"), (4, 8, "adam"), (2, 40, "shared_adam"),
                                             (2, 200, "adam")])
def test_sparse_pull_push_virtual_ranks(nranks, dim, opt):
    """All ranks as virtual ranks of one cooperative launch: the production per-rank code,
    sharded tables, cross-rank mailboxes — checked against a single global oracle."""
    from paddlebox_b200.ops.sparse_table import SparseTable
    from paddlebox_b200.parallel.sparse_engine import EmulatedSparseGroup
    gen = torch.Generator().manual_seed(11)
    B, S = 48, 4
    H = 3 + dim
    tables = [_mk_table(cap=1 << 14, dim=dim, opt=opt) for _ in range(nranks)]
    oracle = _mk_table(cap=1 << 16, dim=dim, opt=opt)
    batches = [_bags(B, S, 5, 2000, gen) for _ in range(nranks)]
    batches = [(k.to(dev()), l.to(dev())) for k, l in batches]
    allk = torch.cat([k for k, _ in batches]).unique()
    # every key lives on the shard pbx_owner_of() picks; prime oracle, then copy rows to owners
    _prime(oracle, allk, dim)
    ok, orow = oracle.dump()

    def mix64(k):
        m = (1 << 64) - 1
        k ^= k >> 33
        k = (k * 0xff51afd7ed558ccd) & m
        k ^= k >> 33
        k = (k * 0xc4ceb9fe1a85ec53) & m
        k ^= k >> 33
        return k
    owner = torch.tensor([(mix64(int(k)) >> 40) % nranks for k in ok.tolist()], device=dev())
    for r in range(nranks):
        sel = owner == r
        tables[r].insert(ok[sel], orow[sel])
    grp = EmulatedSparseGroup(tables, max_keys=max(k.numel() for k, _ in batches) + 8)
    restores = grp.pull([k for k, _ in batches])
    grp.check_status()
    for r, (k, _) in enumerate(batches):
        got = grp.engines[r].values[restores[r].long()][:, :H]
        assert torch.equal(got, oracle.lookup(k)), f"rank {r} pull mismatch"
    # push: owners apply one requester at a time -> oracle applies the ranks' merged pushes in order
    douts = [torch.randn(B, S * H, device=dev()) for _ in range(nranks)]
    cvms = [torch.stack([torch.ones(B, device=dev()), (torch.rand(B, device=dev()) < 0.3).float()], 1)
            for _ in range(nranks)]
    grp.push_pooled(restores, douts, cvms, [l for _, l in batches], B, S, H, 2, 0, -float(B))
    grp.check_status()
    # A key pushed by several requesters is updated once per requester in an order the owner is
    # free to choose (per-row try-lock); Adagrad is not commutative, so for such keys every
    # application order is a valid result (up to 5 pushers: all orders are tried; more: only the
    # order-independent words are checked).
    import itertools
    base = {int(k): r.tolist() for k, r in zip(ok.tolist(), orow.cpu())}
    pushes = {}
    for r, (k, l) in enumerate(batches):
        uk, inv = torch.unique(k, return_inverse=True)
        dx = R.seqpool_cvm_bwd(douts[r].cpu(), cvms[r].cpu(), l.cpu(), B, S, H)
        k2s = torch.repeat_interleave(torch.arange(S), (l[1:] - l[:-1]).cpu().view(S, B).sum(1))
        merged = R.merge_push(inv.cpu(), uk.numel(), dx, k2s, 2, -float(B))
        for i, kk in enumerate(uk.tolist()):
            pushes.setdefault(kk, []).append(merged[i].tolist())
    L, cfg = tables[0].layout, tables[0].config
    slot_col = L.off_tail + 1

    def apply(row, recs):
        for rec in recs:
            row = R.sparse_row_update(row, rec, L, cfg)
        return torch.tensor(row)
    n_multi = 0
    for r in range(nranks):
        dk, dr = tables[r].dump()
        for kk, row in zip(dk.tolist(), dr.cpu()):
            recs = pushes.get(kk, [])
            row = row.clone()
            row[slot_col] = 0
            if len(recs) <= 5:
                ok_any = False
                for perm in itertools.permutations(recs):
                    exp = apply(base[kk], perm)
                    exp[slot_col] = 0
                    if torch.allclose(row, exp, rtol=1e-3, atol=1e-4):
                        ok_any = True
                        break
                assert ok_any, (r, kk, len(recs))
            else:
                exp = apply(base[kk], recs)
                assert torch.allclose(row[:2], exp[:2]) and abs(float(row[L.off_tail] - exp[L.off_tail])) < 1e-3
                assert torch.isfinite(row).all()
            n_multi += len(recs) > 1
    assert n_multi > 0      # the test really exercised keys pushed by several requesters


def test_sparse_mailbox_overflow_is_contained_and_reported():
    """More unique keys for one shard than the mailbox holds: the surplus keys pull zeros and their
    gradients are dropped; no other key's values or table row are affected; status reports it."""
    from paddlebox_b200.parallel.sparse_engine import EmulatedSparseGroup
    gen = torch.Generator().manual_seed(5)
    nranks, dim, B, S = 2, 8, 48, 4
    H = 3 + dim
    tables = [_mk_table(cap=1 << 14, dim=dim) for _ in range(nranks)]
    oracle = _mk_table(cap=1 << 16, dim=dim)
    batches = [_bags(B, S, 5, 2000, gen) for _ in range(nranks)]
    batches = [(k.to(dev()), l.to(dev())) for k, l in batches]
    allk = torch.cat([k for k, _ in batches]).unique()
    _prime(oracle, allk, dim)
    ok, orow = oracle.dump()

    def mix64(k):
        m = (1 << 64) - 1
        k ^= k >> 33
        k = (k * 0xff51afd7ed558ccd) & m
        k ^= k >> 33
        k = (k * 0xc4ceb9fe1a85ec53) & m
        k ^= k >> 33
        return k
    owner = torch.tensor([(mix64(int(k)) >> 40) % nranks for k in ok.tolist()], device=dev())
    for r in range(nranks):
        tables[r].insert(ok[owner == r], orow[owner == r])
    before = [t.dump() for t in tables]
    grp = EmulatedSparseGroup(tables, max_keys=max(k.numel() for k, _ in batches) + 8, cap=64)
    restores = grp.pull([k for k, _ in batches])
    with pytest.raises(RuntimeError, match="overflow"):
        grp.check_status()
    zero_keys = set()
    for r, (k, _) in enumerate(batches):
        got = grp.engines[r].values[restores[r].long()][:, :H]
        exp = oracle.lookup(k)
        same = (got == exp).all(1)
        zero = (got == 0).all(1)
        assert bool((same | zero).all()), "an overflowed key must pull zeros, every other key its own row"
        assert bool(zero.any()) and bool(same.any())
        zero_keys.update(k[zero & ~same].tolist())
    douts = [torch.randn(B, S * H, device=dev()) for _ in range(nranks)]
    cvms = [torch.stack([torch.ones(B, device=dev()), torch.zeros(B, device=dev())], 1) for _ in range(nranks)]
    grp.push_pooled(restores, douts, cvms, [l for _, l in batches], B, S, H, 2, 0, -float(B))
    with pytest.raises(RuntimeError, match="overflow"):
        grp.check_status()
    pushed_ok = set()
    for r, (k, _) in enumerate(batches):
        got = grp.engines[r].values[restores[r].long()][:, :H]
        pushed_ok.update(k[~(got == 0).all(1)].tolist())
    n_untouched = 0
    for r in range(nranks):
        ak, ar = tables[r].dump()
        bmap = {int(kk): row for kk, row in zip(before[r][0].tolist(), before[r][1].cpu())}
        for kk, row in zip(ak.tolist(), ar.cpu()):
            if kk not in pushed_ok:           # overflowed on every requester (or never requested): row untouched
                assert torch.equal(row, bmap[kk]), kk
                n_untouched += 1
            else:
                assert float(row[0]) > float(bmap[kk][0])          # show count grew: the push arrived
    assert n_untouched > 0


# ------------------------------------------------------------------ seqpool
@pytest.mark.parametrize("cfg", [
    dict(),
    dict(use_cvm=False),
    dict(clk_filter=True),
    dict(need_filter=True, threshold=0.5),
    dict(quant_ratio=128),
    dict(need_filter=True, embed_threshold_filter=True, embed_threshold=0.3, embed_thres_size=4, threshold=0.2),
    dict(embedx_concate_size=3),
    dict(pad_value=0.5),
    dict(variant=1, cvm_offset=3),
    dict(variant=2, cvm_offset=4, show_filter=True),
])
def test_seqpool_cvm_fwd_bwd(cfg):
    from paddlebox_b200.ops.seqpool import SeqpoolAttrs, seqpool_cvm_forward, seqpool_cvm_backward
    gen = torch.Generator().manual_seed(2)
    B, S, H = 37, 6, 11
    keys, lod = _bags(B, S, 5, 100, gen)
    n = keys.numel()
    x = torch.rand(n, H, generator=gen)
    x[:, 0] = torch.randint(1, 20, (n,), generator=gen).float()
    x[:, 1] = torch.floor(x[:, 0] * torch.rand(n, generator=gen))
    attrs = SeqpoolAttrs(**cfg)
    exp = R.seqpool_cvm_fwd(x, lod, B, S, **attrs.ref_kwargs())
    got = seqpool_cvm_forward(x.to(dev()), lod.to(dev()), B, S, attrs, H)
    assert got.shape == exp.shape
    assert torch.allclose(got.cpu(), exp, rtol=1e-5, atol=1e-5)
    # through a restore index (values deduped)
    perm = torch.randperm(n, generator=gen)
    xs = torch.zeros(n, 12)
    xs[perm, :H] = x
    got2 = seqpool_cvm_forward(xs.to(dev()), lod.to(dev()), B, S, attrs, H, restore=perm.int().to(dev()))
    assert torch.allclose(got2.cpu(), exp, rtol=1e-5, atol=1e-5)
    dout = torch.randn(B, exp.shape[1], generator=gen)
    cvm = torch.rand(B, attrs.cvm_offset, generator=gen)
    exp_dx = R.seqpool_cvm_bwd(dout, cvm, lod, B, S, H, use_cvm=attrs.use_cvm, cvm_offset=attrs.cvm_offset,
                               clk_filter=attrs.clk_filter, embedx_concate_size=attrs.embedx_concate_size,
                               variant=attrs.variant, show_filter=attrs.show_filter)
    got_dx = seqpool_cvm_backward(dout.to(dev()), cvm.to(dev()), lod.to(dev()), B, S, n, attrs, H)
    assert torch.allclose(got_dx.cpu(), exp_dx)


# ------------------------------------------------------------------ GEMM
@pytest.mark.parametrize("impl", [1, 0])
@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 128, 192), (300, 200, 136), (4096, 512, 496),
                                   (4096, 1, 128), (1, 256, 4096), (77, 33, 40)])
def test_gemm_bf16_tn(impl, M, N, K):
    from paddlebox_b200.ops.mlp import gemm_bf16_tn, _pad8
    torch.manual_seed(M + N + K)
    A = torch.zeros(M, _pad8(K), dtype=torch.bfloat16, device=dev())
    Bm = torch.zeros(N, _pad8(K), dtype=torch.bfloat16, device=dev())
    A[:, :K] = torch.randn(M, K, device=dev()) * 0.5
    Bm[:, :K] = torch.randn(N, K, device=dev()) * 0.5
    bias = torch.randn(N, device=dev())
    ref = A[:, :K].float() @ Bm[:, :K].float().t()
    tol = dict(rtol=2e-2, atol=2e-2 * math.sqrt(K / 64))
    # plain fp32 output
    out = torch.full((M, N), float("nan"), device=dev())
    gemm_bf16_tn(A, Bm, M, N, K, out_f32=out, impl=impl)
    assert torch.allclose(out, ref, **tol), (out - ref).abs().max()
    # bias + relu, bf16 + transposed outputs
    ob = torch.zeros(M, _pad8(N), dtype=torch.bfloat16, device=dev())
    ot = torch.zeros(N, _pad8(M), dtype=torch.bfloat16, device=dev())
    gemm_bf16_tn(A, Bm, M, N, K, bias=bias, out_bf16=ob, out_t=ot, relu=True, impl=impl)
    r2 = torch.relu(ref + bias)
    assert torch.allclose(ob[:, :N].float(), r2, rtol=3e-2, atol=tol["atol"] + 2e-2)
    assert torch.equal(ot[:, :M], ob[:, :N].t())
    # mask epilogue (ReLU backward) and alpha
    mask = torch.zeros(M, _pad8(N), dtype=torch.bfloat16, device=dev())
    mask[:, :N] = torch.randn(M, N, device=dev())
    om = torch.full((M, N), float("nan"), device=dev())
    gemm_bf16_tn(A, Bm, M, N, K, mask=mask, out_f32=om, alpha=0.5, impl=impl)
    assert torch.allclose(om, 0.5 * ref * (mask[:, :N].float() > 0), **tol)
    # split-K accumulate
    acc = torch.ones(M, N, device=dev())
    gemm_bf16_tn(A, Bm, M, N, K, out_f32=acc, accumulate=True, splitk=4 if impl == 0 else 1, impl=impl)
    assert torch.allclose(acc, ref + 1, **tol)
    # bias-gradient column: the last output column is routed to a separate vector
    if N > 1:
        acc2 = torch.zeros(M, N - 1, device=dev())
        colv = torch.zeros(M, device=dev())
        gemm_bf16_tn(A, Bm, M, N, K, out_f32=acc2, accumulate=True, splitk=2 if impl == 0 else 1, impl=impl,
                     col_out=colv, col_split=N - 1)
        assert torch.allclose(acc2, ref[:, :N - 1], **tol) and torch.allclose(colv, ref[:, N - 1], **tol)


def test_fused_mlp_matches_torch():
    from paddlebox_b200.ops.mlp import FusedMLP
    torch.manual_seed(0)
    dims = [200, 128, 64, 1]
    m = FusedMLP(dims, seed=1).to(dev())
    x = torch.randn(512, dims[0], device=dev(), requires_grad=True)
    y = m(x)
    # torch reference that rounds to bf16 at the same points as the kernels (operands and stored
    # activations), so that the ReLU masks agree; accumulation stays fp32
    def rb(t):
        return t.to(torch.bfloat16).float()
    h = x.detach().clone().requires_grad_(True)
    ws = [w.detach().clone().requires_grad_(True) for w in m.weights]
    bs = [b.detach().clone().requires_grad_(True) for b in m.biases]

    class _RoundSTE(torch.autograd.Function):
        @staticmethod
        def forward(ctx, t):
            return rb(t)

        @staticmethod
        def backward(ctx, g):
            return rb(g)
    z = _RoundSTE.apply(h)
    for i, (w, b) in enumerate(zip(ws, bs)):
        z = torch.nn.functional.linear(z, _RoundSTE.apply(w), b)
        if i < len(ws) - 1:
            z = _RoundSTE.apply(torch.relu(z))
    assert torch.allclose(y, z, rtol=1e-2, atol=1e-2), (y - z).abs().max()
    g = torch.randn_like(y)
    y.backward(g)
    z.backward(g)
    assert torch.allclose(x.grad, h.grad, rtol=3e-2, atol=3e-2), (x.grad - h.grad).abs().max()
    for w, wr in zip(m.weights, ws):
        scale = wr.grad.abs().max().item()
        assert torch.allclose(w.grad, wr.grad, rtol=3e-2, atol=3e-2 * max(scale, 1)), (w.grad - wr.grad).abs().max()
    for b, br in zip(m.biases, bs):
        assert torch.allclose(b.grad, br.grad, rtol=3e-2, atol=3e-2 * max(br.grad.abs().max().item(), 1))


# ------------------------------------------------------------------ dense helpers
def test_dense_helper_kernels():
    from paddlebox_b200.ops import functional as F
    torch.manual_seed(0)
    # cvm
    x = torch.rand(100, 9, device=dev()) * 5
    cvm = torch.rand(100, 2, device=dev())
    for use in (True, False):
        xr = x.clone().requires_grad_(True)
        y = F.continuous_value_model(xr, cvm, use)
        assert torch.allclose(y, R.cvm_fwd(x, use), atol=1e-6)
        g = torch.randn_like(y)
        y.backward(g)
        assert torch.allclose(xr.grad, R.cvm_bwd(g, cvm, 9, use))
    # data_norm
    xs = torch.randn(300, 40, device=dev()) * 3 + 1
    bs, bsum, bsq = torch.full((40,), 1e4, device=dev()), torch.randn(40, device=dev()) * 1e3, \
        torch.full((40,), 2e4, device=dev())
    got = {}
    xr = xs.clone().requires_grad_(True)
    y = F.data_norm(xr, bs, bsum, bsq, 1e-4, lambda a, b, c: got.update(s=a.clone(), m=b.clone(), q=c.clone()))
    ey, mean, scale = R.data_norm_fwd(xs, bs, bsum, bsq)
    assert torch.allclose(y, ey, rtol=1e-5, atol=1e-5)
    g = torch.randn_like(y)
    y.backward(g)
    edx, es, em, eq = R.data_norm_bwd(xs, g, mean, scale, 1e-4)
    assert torch.allclose(xr.grad, edx, rtol=1e-5, atol=1e-6)
    assert torch.allclose(got["s"], es) and torch.allclose(got["m"], em, rtol=1e-4, atol=1e-4)
    assert torch.allclose(got["q"], eq, rtol=1e-4, atol=1e-3)
    # fm
    S, E, D = 7, 11, 8
    xp = (torch.randn(64, S * E + 5, device=dev()) * 0.3).requires_grad_(True)
    fm = F.fm_interaction(xp, S, E, 2, 3, D)
    xq = xp.detach().clone().requires_grad_(True)
    efm = R.fm_fwd(xq, S, E, 2, 3, D)
    assert torch.allclose(fm, efm, rtol=1e-4, atol=1e-5)
    g = torch.randn(64, device=dev())
    fm.backward(g)
    efm.backward(g)
    assert torch.allclose(xp.grad, xq.grad, rtol=1e-4, atol=1e-5)
    # sigmoid + logloss
    z = torch.randn(1000, device=dev(), requires_grad=True)
    yl = (torch.rand(1000, device=dev()) < 0.3).float()
    loss, pred = F.sigmoid_log_loss(z, yl)
    z2 = z.detach().clone().requires_grad_(True)
    el = torch.nn.functional.binary_cross_entropy_with_logits(z2, yl)
    assert torch.allclose(loss, el, rtol=1e-5) and torch.allclose(pred, torch.sigmoid(z2), atol=1e-6)
    loss.backward()
    el.backward()
    assert torch.allclose(z.grad, z2.grad, atol=1e-7)
    # adam flat vs torch.optim.Adam
    p = torch.randn(10000, device=dev())
    pt = p.clone().requires_grad_(True)
    opt = torch.optim.Adam([pt], lr=1e-3)
    m, v, step = torch.zeros_like(p), torch.zeros_like(p), torch.zeros(2, device=dev())
    for _ in range(3):
        g = torch.randn_like(p)
        pt.grad = g.clone()
        opt.step()
        N.check(N.cuda().pbx_adam_flat(N.ptr(p), N.ptr(g), N.ptr(m), N.ptr(v), C.c_int64(p.numel()), C.c_float(1e-3),
                                       C.c_float(0.9), C.c_float(0.999), C.c_float(1e-8), C.c_float(0.0),
                                       C.c_float(1.0), N.ptr(step), 1, N.stream_ptr()))
        assert torch.all(g == 0)
    assert torch.allclose(p, pt.detach(), rtol=1e-5, atol=1e-6)
    # row gather
    tab = torch.randn(50, 8, device=dev())
    idx = torch.randint(0, 50, (200,), device=dev())
    assert torch.equal(F.row_gather(tab, idx), tab[idx])


def test_auc_bucket_kernel():
    n, nb = 20000, 1000
    pred = torch.rand(n, device=dev())
    label = (torch.rand(n, device=dev()) < pred).float()
    table = torch.zeros(2, nb, dtype=torch.float64, device=dev())
    sums = torch.zeros(5, dtype=torch.float64, device=dev())
    N.check(N.cuda().pbx_auc_bucket(N.ptr(pred), N.ptr(label), None, C.c_int64(n), nb, N.ptr(table), N.ptr(sums),
                                    N.stream_ptr()))
    bk = torch.clamp((pred * nb).long(), max=nb - 1)
    exp = torch.zeros(2, nb, dtype=torch.float64, device=dev())
    exp[0].index_add_(0, bk, (1 - label).double())
    exp[1].index_add_(0, bk, label.double())
    assert torch.equal(table, exp)
    assert abs(sums[4].item() - n) < 1e-9
    assert abs(sums[0].item() - (pred - label).abs().double().sum().item()) < 1e-3


def test_batch_build_kernel():
    gen = torch.Generator().manual_seed(4)
    B, S = 300, 9
    lens = torch.randint(0, 4, (B, S), generator=gen)
    rec_off = torch.zeros(B, S + 1, dtype=torch.int32)
    total = 0
    for b in range(B):
        rec_off[b, 0] = total
        rec_off[b, 1:] = total + torch.cumsum(lens[b], 0)
        total = int(rec_off[b, S])
    rec_keys = torch.randint(1, 1 << 60, (total,), generator=gen, dtype=torch.int64)
    keys_out = torch.zeros(total, dtype=torch.int64, device=dev())
    lod = torch.zeros(S * B + 1, dtype=torch.int32, device=dev())
    scratch = torch.zeros(1024, dtype=torch.int32, device=dev())
    d_keys, d_off = rec_keys.to(dev()), rec_off.to(dev())
    N.check(N.cuda().pbx_batch_build(N.ptr(d_keys), N.ptr(d_off), B, S, N.ptr(keys_out),
                                     N.ptr(lod), N.ptr(scratch), N.stream_ptr()))
    torch.cuda.synchronize()
    exp_lens = lens.t().reshape(-1)
    exp_lod = torch.zeros(S * B + 1, dtype=torch.int32)
    exp_lod[1:] = torch.cumsum(exp_lens, 0)
    assert torch.equal(lod.cpu(), exp_lod)
    exp_keys = torch.cat([rec_keys[rec_off[b, s]:rec_off[b, s + 1]] for s in range(S) for b in range(B)])
    assert torch.equal(keys_out.cpu(), exp_keys)


@pytest.mark.parametrize("nranks", [2, 8])
def test_symm_collectives_virtual_ranks(nranks):
    from paddlebox_b200.parallel.symm import SymmetricBuffer
    n = 10007
    bufs = SymmetricBuffer.emulated_group(nranks, n)
    data = [torch.randn(n, device=dev()) for _ in range(nranks)]
    for b, d in zip(bufs, data):
        b.tensor[:n].copy_(d)
    SymmetricBuffer.collective_emu(bufs, 0, n, 1.0 / nranks, "allreduce")
    exp = torch.stack(data).sum(0) / nranks
    for b in bufs:
        assert torch.allclose(b.tensor[:n], exp, rtol=1e-5, atol=1e-6)
    # broadcast
    bufs[1].tensor[:n].copy_(data[0])
    SymmetricBuffer.collective_emu(bufs, 0, n, 1.0, "broadcast", root=1)
    for b in bufs:
        assert torch.equal(b.tensor[:n], data[0])


def test_ctr_head_matches_torch():
    from paddlebox_b200.ops import functional as F
    torch.manual_seed(0)
    B = 1000
    tower = torch.randn(B, 1, device=dev(), requires_grad=True)
    extra = torch.randn(B, device=dev(), requires_grad=True)
    bias = torch.randn(1, device=dev(), requires_grad=True)
    dense = torch.rand(B, 5, device=dev())
    dense[:, 0] = (dense[:, 0] < 0.3).float()
    label = dense[:, 0]                      # strided column, as the models pass it
    loss, pred = F.ctr_head(tower, label, extra=extra, bias=bias)
    (loss * 2.0).backward()
    t2, e2, b2 = [x.detach().clone().requires_grad_(True) for x in (tower, extra, bias)]
    z = t2.view(-1) + e2 + b2
    ref = torch.nn.functional.binary_cross_entropy_with_logits(z, label)
    (ref * 2.0).backward()
    assert torch.allclose(loss, ref, rtol=1e-5, atol=1e-6)
    assert torch.allclose(pred, torch.sigmoid(z), atol=1e-6)
    assert torch.allclose(tower.grad, t2.grad, atol=1e-7) and torch.allclose(extra.grad, e2.grad, atol=1e-7)
    assert torch.allclose(bias.grad, b2.grad, rtol=1e-4, atol=1e-6)
    # tower only
    t3 = tower.detach().clone().requires_grad_(True)
    l3, _ = F.ctr_head(t3, label)
    l3.backward()
    t4 = tower.detach().clone().requires_grad_(True)
    torch.nn.functional.binary_cross_entropy_with_logits(t4.view(-1), label).backward()
    assert torch.allclose(t3.grad, t4.grad, atol=1e-7)


def test_cast_transpose_multi():
    from paddlebox_b200.ops.mlp import cast_transpose_multi, _pad8
    torch.manual_seed(1)
    shapes = [(512, 507), (256, 512), (1, 128), (77, 33)] + [(40, 9)] * 10      # > 12 tensors: two launches
    items, refs = [], []
    for i, (M, K) in enumerate(shapes):
        x = torch.randn(M, K, device=dev())
        out = torch.zeros(M, _pad8(K), dtype=torch.bfloat16, device=dev()) if i != 2 else None
        out_t = torch.zeros(K, _pad8(M), dtype=torch.bfloat16, device=dev()) if i != 3 else None
        items.append((x, out, out_t))
    cast_transpose_multi(items)
    for (x, out, out_t) in items:
        M, K = x.shape
        xb = x.to(torch.bfloat16)
        if out is not None:
            assert torch.equal(out[:, :K], xb) and float(out[:, K:].abs().sum()) == 0
        if out_t is not None:
            assert torch.equal(out_t[:, :M], xb.t())
