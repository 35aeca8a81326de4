"""Ragged-sequence operators and the ranking / distillation losses against plain per-sequence loops."""
import math

import numpy as np
import pytest
import torch

from paddlebox_b200 import layers as L
from paddlebox_b200.ops import sequence as S


def _ragged(n_seq=7, d=5, seed=0, empty=(2,), max_len=6):
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(1, max_len, (n_seq,), generator=g)
    for e in empty:
        lens[e] = 0
    lod = torch.zeros(n_seq + 1, dtype=torch.int64)
    lod[1:] = lens.cumsum(0)
    x = torch.randn(int(lod[-1]), d, generator=g)
    return x, lod


def _seqs(x, lod):
    lod = [int(v) for v in lod]
    return [x[a:b] for a, b in zip(lod[:-1], lod[1:])]


@pytest.mark.parametrize("kind", ["sum", "average", "sqrt", "max", "min", "first", "last"])
def test_sequence_pool_matches_a_loop_and_pads_empty_sequences(kind):
    x, lod = _ragged()
    x.requires_grad_(True)
    got = S.sequence_pool(x, lod, kind, pad_value=-7.0)
    fn = {"sum": lambda s: s.sum(0), "average": lambda s: s.mean(0), "sqrt": lambda s: s.sum(0) / math.sqrt(len(s)),
          "max": lambda s: s.max(0).values, "min": lambda s: s.min(0).values, "first": lambda s: s[0], "last": lambda s: s[-1]}[kind]
    want = torch.stack([fn(s) if len(s) else torch.full((x.shape[1],), -7.0) for s in _seqs(x, lod)])
    torch.testing.assert_close(got, want)
    g1, = torch.autograd.grad(got.sum(), x, retain_graph=True)
    g2, = torch.autograd.grad(want.sum(), x)
    torch.testing.assert_close(g1, g2)


def test_first_last_step_and_softmax():
    x, lod = _ragged(empty=())
    torch.testing.assert_close(S.sequence_first_step(x, lod), torch.stack([s[0] for s in _seqs(x, lod)]))
    torch.testing.assert_close(S.sequence_last_step(x, lod), torch.stack([s[-1] for s in _seqs(x, lod)]))
    v = x[:, :1].clone().requires_grad_(True)
    got = S.sequence_softmax(v, lod)
    want = torch.cat([torch.softmax(s, 0) for s in _seqs(v, lod)])
    torch.testing.assert_close(got, want)
    w = torch.randn_like(got)
    torch.testing.assert_close(torch.autograd.grad((got * w).sum(), v, retain_graph=True)[0],
                               torch.autograd.grad((want * w).sum(), v)[0])


def test_concat_expand_pad_unpad_reverse_slice():
    a, la = _ragged(seed=1)
    b, lb = _ragged(seed=2, empty=(4,))
    out, lo = S.sequence_concat([a, b], [la, lb])
    want = [torch.cat([p, q]) for p, q in zip(_seqs(a, la), _seqs(b, lb))]
    torch.testing.assert_close(out, torch.cat(want))
    assert lo.tolist() == [0] + list(np.cumsum([len(w) for w in want]))

    rows = torch.randn(la.numel() - 1, 3)
    torch.testing.assert_close(S.sequence_expand_as(rows, la), torch.cat([rows[i:i + 1].expand(len(s), 3) for i, s in enumerate(_seqs(a, la))]))
    torch.testing.assert_close(S.sequence_expand(rows, la), S.sequence_expand_as(rows, la))
    out, lo = S.sequence_expand(a, lb, x_lod=la)                     # sequence i of a repeated len_b[i] times
    want, lens = [], []
    for s, t in zip(_seqs(a, la), _seqs(b, lb)):
        for _ in range(len(t)):
            want.append(s)
            lens.append(len(s))
    torch.testing.assert_close(out, torch.cat(want))
    assert lo.tolist() == [0] + list(np.cumsum(lens))

    padded, lens = S.sequence_pad(a, la, pad_value=9.0)
    assert padded.shape == (la.numel() - 1, int(lens.max()), a.shape[1])
    for i, s in enumerate(_seqs(a, la)):
        torch.testing.assert_close(padded[i, :len(s)], s)
        assert bool((padded[i, len(s):] == 9.0).all())
    back, lod2 = S.sequence_unpad(padded, lens)
    torch.testing.assert_close(back, a)
    assert lod2.tolist() == la.tolist()
    with pytest.raises(ValueError):
        S.sequence_pad(a, la, maxlen=1)

    torch.testing.assert_close(S.sequence_reverse(a, la), torch.cat([s.flip(0) for s in _seqs(a, la)]))

    lens = la[1:] - la[:-1]
    off = (lens > 1).long()
    ln = (lens - off).clamp(max=2)
    out, lo = S.sequence_slice(a, la, off.view(-1, 1), ln.view(-1, 1))
    torch.testing.assert_close(out, torch.cat([s[o:o + n] for s, o, n in zip(_seqs(a, la), off.tolist(), ln.tolist())]))
    assert lo.tolist() == [0] + ln.cumsum(0).tolist()
    with pytest.raises(ValueError):
        S.sequence_slice(a, la, off, lens + 1)


def test_mask_enumerate_reshape_scatter():
    assert S.sequence_mask(torch.tensor([3, 1, 0]), maxlen=4).tolist() == [[1, 1, 1, 0], [1, 0, 0, 0], [0, 0, 0, 0]]
    assert S.sequence_mask(torch.tensor([2, 3]), dtype=torch.float32).shape == (2, 3)
    ids = torch.tensor([1, 2, 3, 4, 5, 6])
    out = S.sequence_enumerate(ids, [0, 3, 3, 6], win_size=2, pad_value=0)
    assert out.tolist() == [[1, 2], [2, 3], [3, 0], [4, 5], [5, 6], [6, 0]]
    x = torch.arange(24.0).view(6, 4)
    out, lod = S.sequence_reshape(x, [0, 2, 6], 8)
    assert out.shape == (3, 8) and lod.tolist() == [0, 1, 3]
    with pytest.raises(ValueError):
        S.sequence_reshape(x, [0, 1, 6], 8)
    base = torch.ones(3, 6)
    out = S.sequence_scatter(base, torch.tensor([[1], [1], [5], [0], [2], [3]]), torch.tensor([.3, .3, .4, .1, .2, .5]),
                             [0, 3, 4, 6])
    want = torch.ones(3, 6)
    want[0, 1] += .6; want[0, 5] += .4; want[1, 0] += .1; want[2, 2] += .2; want[2, 3] += .5   # noqa: E702
    torch.testing.assert_close(out, want)


def test_sequence_conv_is_a_zero_padded_context_window():
    x, lod = _ragged(d=3, empty=(1,))
    w = torch.randn(3 * 3, 4)
    bias = torch.randn(4)
    got = S.sequence_conv(x, lod, w, bias, filter_size=3, act="tanh")
    want = []
    for s in _seqs(x, lod):
        p = torch.cat([torch.zeros(1, 3), s, torch.zeros(1, 3)])
        for t in range(len(s)):
            want.append(torch.tanh(p[t:t + 3].reshape(-1) @ w + bias))
    torch.testing.assert_close(got, torch.stack(want))
    got = S.sequence_conv(x, lod, torch.randn(2 * 3, 4), filter_size=2, padding_start=0)   # looks ahead only
    assert got.shape == (x.shape[0], 4)


def test_match_matrix_and_topk_avg_pooling():
    x, lx = _ragged(n_seq=4, d=6, seed=3, empty=())
    y, ly = _ragged(n_seq=4, d=6, seed=4, empty=())
    c = 3
    w = torch.randn(6, c, 6)
    out, lo, tmp = S.match_matrix_tensor(x, lx, y, ly, w)
    want, sizes = [], []
    for a, b in zip(_seqs(x, lx), _seqs(y, ly)):
        m = torch.stack([a @ w[:, k, :] @ b.t() for k in range(c)])
        want.append(m.reshape(-1))
        sizes.append(m.numel())
    torch.testing.assert_close(out.reshape(-1), torch.cat(want), rtol=1e-4, atol=1e-4)
    assert lo.tolist() == [0] + list(np.cumsum(sizes)) and tmp.shape == (x.shape[0], c * 6)

    topks = [1, 3, 7]
    got, rl = S.sequence_topk_avg_pooling(out, lo, lx, ly, topks, c)
    rows = []
    for a, b, m in zip(_seqs(x, lx), _seqs(y, ly), want):
        m = m.view(c, len(a), len(b))
        for r in range(len(a)):
            rows.append(torch.stack([m[k, r].sort(descending=True).values[:t].sum() / t for k in range(c) for t in topks]))
    torch.testing.assert_close(got, torch.stack(rows), rtol=1e-4, atol=1e-4)
    assert rl.tolist() == lx.tolist()


def test_ranking_losses():
    g = torch.Generator().manual_seed(0)
    left, right = torch.randn(9, 1, generator=g), torch.randn(9, 1, generator=g)
    label = (torch.rand(9, 1, generator=g) > 0.5).float()
    o = left - right
    torch.testing.assert_close(S.rank_loss(label, left, right), torch.log(1 + torch.exp(o)) - label * o)
    pm = label * 2 - 1
    torch.testing.assert_close(S.margin_rank_loss(pm, left, right, 0.2), torch.clamp(-pm * o + 0.2, min=0))
    x = torch.randn(5, 4, generator=g)
    y = torch.tensor([[0], [3], [1], [2], [3]])
    want = [-sum(math.log(1 / (1 + math.exp(-(float(x[i, int(y[i])]) - float(x[i, j])))))
                 for j in range(4) if j != int(y[i])) / 3 for i in range(5)]
    torch.testing.assert_close(S.bpr_loss(x, y).reshape(-1), torch.tensor(want), rtol=1e-5, atol=1e-6)
    p = torch.rand(6, 1, generator=g)
    lab = (torch.rand(6, 1, generator=g) > 0.5).float()
    torch.testing.assert_close(S.log_loss(p, lab), -lab * torch.log(p + 1e-4) - (1 - lab) * torch.log(1 - p + 1e-4))


def test_teacher_student_loss_value_and_bounded_gradient():
    x = torch.tensor([0.3, -1.2, 2.0, 0.7, 20.0, -20.0], requires_grad=True)
    label = torch.tensor([-2.0, -1.0, 0.25, 1.6, 0.5, -1.0])
    loss = S.teacher_student_sigmoid_loss(x, label)
    ce = lambda v, z: max(v, 0) - v * z + math.log(1 + math.exp(-abs(v)))   # noqa: E731
    want = [ce(0.3, 0), ce(-1.2, 1), ce(2.0, 0) + ce(2.0, 0.25), ce(0.7, 1) + ce(0.7, 0.6), ce(20.0, 0) + ce(20.0, 0.5), ce(-20.0, 1)]
    torch.testing.assert_close(loss, torch.tensor(want), rtol=1e-5, atol=1e-6)
    loss.sum().backward()
    sg = lambda v: 1 / (1 + math.exp(-v))                                  # noqa: E731
    want_g = [sg(0.3), sg(-1.2) - 1, 2 * sg(2.0) - 0.25, 2 * sg(0.7) - 1.6, 0.0, 0.0]   # zero outside the bounds
    torch.testing.assert_close(x.grad, torch.tensor(want_g), rtol=1e-5, atol=1e-6)


def test_hash_is_xxh64_of_the_row_bytes():
    xxhash = pytest.importorskip("xxhash")
    ids = torch.tensor([[1, 2], [3, 4], [-5, 2 ** 40]], dtype=torch.int64)
    out = L.hash(ids, hash_size=10007, num_hash=3)
    assert out.shape == (3, 3, 1) and out.dtype == torch.int64
    for i in range(3):
        for k in range(3):
            assert int(out[i, k, 0]) == xxhash.xxh64(ids[i].numpy().tobytes(), seed=k).intdigest() % 10007
    wide = torch.arange(3 * 9, dtype=torch.int32).view(3, 9)              # 36 bytes: the 32-byte stripe loop
    out = L.hash(wide, hash_size=1 << 20)
    assert [int(v) for v in out.view(-1)] == [xxhash.xxh64(wide[i].numpy().tobytes(), seed=0).intdigest() % (1 << 20) for i in range(3)]
    with pytest.raises(TypeError):
        L.hash(torch.zeros(2, 2), 10)


def test_slot_lod_feeds_the_ragged_ops():
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch
    lod = torch.tensor([0, 2, 3, 3, 6, 7, 9], dtype=torch.int32)          # 2 slots x 3 instances
    b = SlotBatch(keys=torch.arange(9), lod=lod, batch_size=3, num_slots=2)
    assert b.slot_lod(0).tolist() == [0, 2, 3, 3] and b.slot_lod(1).tolist() == [0, 3, 4, 6]
    emb = torch.arange(9.0).view(9, 1)
    (b0, e0), (b1, e1) = b.slot_ranges()
    assert S.sequence_pool(emb[b1:e1], b.slot_lod(1), "sum").view(-1).tolist() == [3 + 4 + 5, 6, 7 + 8]


def test_auc_layer_statistics_feed_fleet_metrics():
    from paddlebox_b200 import fleet
    from paddlebox_b200.host import HostAuc
    g = torch.Generator().manual_seed(2)
    layer = L.Auc(num_thresholds=4095, slide_steps=2)
    preds, labels = [], []
    for step in range(4):
        p = torch.rand(500, generator=g)
        y = (torch.rand(500, generator=g) < p).long()
        preds.append(p), labels.append(y)
        total, batch, stats = layer(torch.stack([1 - p, p], 1), y.view(-1, 1))
    P, Y = torch.cat(preds), torch.cat(labels)
    # exact AUC by ranks for comparison (bucketing costs a little precision)
    order = P.argsort()
    ranks = torch.empty_like(order, dtype=torch.double)
    ranks[order] = torch.arange(1, P.numel() + 1, dtype=torch.double)
    npos, nneg = int(Y.sum()), int((1 - Y).sum())
    exact = (float(ranks[Y == 1].sum()) - npos * (npos + 1) / 2) / (npos * nneg)
    assert abs(total - exact) < 2e-3
    last2 = L.Auc.area(*[sum(s) for s in zip(*[(torch.bincount((p * 4095).long()[y == 1], minlength=4096),
                                                torch.bincount((p * 4095).long()[y == 0], minlength=4096))
                                               for p, y in zip(preds[-2:], labels[-2:])])])
    assert abs(batch - last2) < 1e-12 and int(stats[2].sum() + stats[3].sum()) == 2000 and int(stats[0].sum() + stats[1].sum()) == 1000
    assert abs(fleet.metrics.auc(stats[2], stats[3]) - total) < 1e-12          # one worker: the same number
    a1 = L.auc(P, Y, name="m1")[0]
    a2 = L.auc(P, Y, name="m1")[0]                                            # same name: statistics accumulate
    assert abs(a1 - a2) < 1e-12 and int(L.auc(P, Y, name="m2")[2][2].sum()) == npos
    layer.reset()
    assert int(layer.stat_pos.sum()) == 0
