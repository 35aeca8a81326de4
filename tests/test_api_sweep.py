"""Calls the public entry points that no other test reaches, with a semantic check where one is cheap:
rarely-used API surface is where latent crashes hide (Dataset.stats() used to be one)."""
import io
import os

import numpy as np
import pytest
import torch

from paddlebox_b200.box_wrapper import BoxWrapper
from paddlebox_b200.dataset import DatasetFactory
from tests.util_data import slot_vars, write_files


@pytest.fixture
def box():
    BoxWrapper.reset_instance()
    b = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    b.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
    yield b
    BoxWrapper.reset_instance()


def test_dataset_setters_disk_mode_and_sampling(tmp_path, box):
    files = write_files(str(tmp_path), n_files=2, lines_per_file=200, n_sparse=3, n_dense=2, key_space=30)
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.init(batch_size=16, thread_num=2, use_var=slot_vars(3, 2), pipe_command="cat")
    for name, arg in (("set_parse_content", True), ("set_merge_by_sid", True), ("set_queue_num", 4),
                      ("set_fleet_send_batch_size", 512), ("set_fleet_send_sleep_seconds", 1), ("set_pv_batch_size", 8),
                      ("set_ads_offset", "ads"), ("set_rank_offset", "rank_offset"), ("set_shuffle_seed", 3),
                      ("set_max_rank", 3), ("set_trainer_num", 1)):
        getattr(ds, name)(arg)
    ds.set_fea_eval(100, True)
    ds.disable_polling(False)
    ds.disable_shuffle(False)
    assert ds.num_sparse_slots == 3 and ds.dense_dim == 3
    ds.set_filelist(files)
    ds.set_date("20240101")
    # Bernoulli sampling keeps roughly the asked share of the lines
    ds.set_sample_rate(0.5)
    ds.read_ins_into_memory()
    n_half = ds.get_memory_data_size()
    assert 120 < n_half < 280 and ds.get_shuffle_data_size() == n_half == ds.get_pv_data_size()
    ds.release_memory()
    ds.set_sample_rate(1.0)
    # disk mode: parse once into an archive, load the archive later
    ds.set_archivefile(str(tmp_path / "pass.bin"))
    ds.load_into_disk()
    assert ds.get_memory_data_size() == 0 and os.path.getsize(tmp_path / "pass.bin") > 0
    ds.wait_load_disk_done()
    assert ds.get_memory_data_size() == 400
    ds.dynamic_adjust_readers_num(2)
    assert ds._prepared_workers == 2
    ds.preload_into_memory()
    ds.wait_feed_pass_done()
    assert "records=" in ds.stats() and "lines=" in ds.stats()


def test_box_wrapper_phase_metric_registry_and_merge(tmp_path, box):
    box.init_metric("AucCalculator", "join", "label", "pred", metric_phase=1, bucket_size=1000)
    box.init_metric("AucCalculator", "update", "label", "pred", metric_phase=0, bucket_size=1000)
    box.init_metric("NanInfCalculator", "nan", "label", "pred", metric_phase=1, bucket_size=1000)
    assert box.get_metric_name_list(1) == ["join", "nan"] and box.get_metric_name_list(0) == ["update"]
    assert set(box.get_metric_name_list()) == {"join", "update", "nan"}
    assert box.phase() == 1
    box.add_metric_data({"pred": torch.tensor([0.2, float("nan"), float("inf"), 0.9]), "label": torch.tensor([0., 1., 0., 1.])})
    box.flip_phase()
    assert box.phase() == 0
    box.add_metric_data({"pred": torch.tensor([0.3, 0.6]), "label": torch.tensor([0., 1.])})
    assert box.get_metric_msg("update")[7] == 2 and box.get_metric_msg("join")[7] == 4
    nan_msg = box.get_nan_inf_metric_msg("nan")
    assert nan_msg[0] == 1 and nan_msg[1] == 1
    box.set_phase(1)
    # merge_model: rows of another checkpoint replace / extend the table
    keys = np.arange(1, 11, dtype=np.uint64)
    rows = box.host_table.fetch_rows(keys)
    rows[:, 2] = 7.0
    box.host_table.store_rows(keys, rows)
    box.save_base(str(tmp_path / "m"))
    rows[:, 2] = 1.0
    box.host_table.store_rows(keys[:5], rows[:5])
    more = np.arange(100, 103, dtype=np.uint64)
    box.host_table.fetch_rows(more)
    assert box.merge_model(str(tmp_path / "m")) == 10
    assert np.all(box.host_table.fetch_rows(keys, create=False)[:, 2] == 7.0) and box.host_table.size() == 13
    box.shrink_resource()


def test_seqpool_layer_variants_against_the_loop_oracle():
    from paddlebox_b200.ops import reference as R
    from paddlebox_b200.ops import seqpool as SP
    gen = torch.Generator().manual_seed(0)
    B, S, H = 6, 3, 9
    lens = torch.randint(1, 4, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    n = int(lod[-1])
    x = (torch.rand(n, H, generator=gen) * 3).requires_grad_(True)
    cvm = torch.rand(B, 4, generator=gen)
    cases = [(SP.fused_seqpool_cvm_with_conv, {}, dict(variant=1, cvm_offset=3)),
             (SP.fused_seqpool_cvm_with_conv, dict(show_filter=True), dict(variant=1, cvm_offset=3, show_filter=True)),
             (SP.fused_seqpool_cvm_with_credit, {}, dict(variant=2, cvm_offset=4))]
    for fn, kw, ref_kw in cases:
        y = fn(x, cvm, lod, B, S, **kw)
        assert torch.allclose(y, R.seqpool_cvm_fwd(x.detach(), lod, B, S, **ref_kw), atol=1e-5)
        y.sum().backward()
        co = ref_kw["cvm_offset"]
        b_of = torch.repeat_interleave(torch.arange(S * B), lens) % B
        assert torch.allclose(x.grad[:, :co], cvm[b_of, :co]) and torch.all(x.grad[:, co:] == 1)
        x.grad = None
    thr = [0.0, 1e9, 0.5]                                           # slot 1 filters everything out
    y = SP.fused_seqpool_cvm_with_diff_thres(x.detach(), cvm, lod, B, S, thr, use_cvm=False)
    ref = R.seqpool_cvm_fwd(x.detach(), lod, B, S, use_cvm=False, need_filter=True, quant_ratio=128,
                            slot_threshold=torch.tensor(thr))
    assert torch.allclose(y, ref, atol=1e-5) and torch.all(y[:, (H - 2):2 * (H - 2)] == 0)
    tw = torch.rand(n, generator=gen)
    y = SP.fused_seqpool_cvm_tradew(x.detach(), cvm, lod, B, S, tw, use_cvm=False)
    want = torch.zeros(S * B, H - 2).index_add_(0, torch.repeat_interleave(torch.arange(S * B), lens),
                                                x.detach()[:, 2:] * tw[:, None])
    assert torch.allclose(y, want.view(S, B, -1).permute(1, 0, 2).reshape(B, -1), atol=1e-5)
    parts = SP.split_slots(y, S)
    assert len(parts) == S and parts[1].shape == (B, H - 2)


def test_extended_sparse_rank_attention2_seq_tensor_dedup(box):
    from paddlebox_b200.ops import ctr_ops as CO
    from paddlebox_b200.ops import reference as R
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch, pull_box_extended_sparse
    # expand embedding = trailing embedx columns (dim 4 table: size 3 + 2, extend 2)
    keys = torch.tensor([5, 6, 5, 7], dtype=torch.int64)
    box.begin_feed_pass("20240101"); box.feed_pass(keys); box.end_feed_pass(); box.begin_pass()
    hk = np.array([5, 6, 7], np.uint64)
    rows = box.host_table.fetch_rows(hk)
    rows[:, 3:7] = np.arange(12, dtype=np.float32).reshape(3, 4)
    rows[:, box.host_table.layout.off_mf] = 4
    box.host_table.store_rows(hk, rows)
    batch = SlotBatch(keys, torch.arange(5, dtype=torch.int32), 4, 1)
    emb, ext = pull_box_extended_sparse(batch, size=5, extend_size=2, box=box)
    assert emb.shape == (4, 5) and ext.shape == (4, 2)
    assert ext[0].tolist() == [2.0, 3.0] and ext[1].tolist() == [6.0, 7.0] and torch.equal(ext[0], ext[2])
    with pytest.raises(ValueError):
        pull_box_extended_sparse(batch, size=5, extend_size=3, box=box)
    # expand_only=False ("WithEmb" kernel): the extended output carries the base row too; offset drops cvm columns
    emb2, ext2 = pull_box_extended_sparse(batch, size=5, extend_size=2, box=box, offset=2, expand_only=False)
    assert emb2.shape == (4, 3) and ext2.shape == (4, 5)
    assert torch.equal(ext2[:, :3], emb2) and torch.equal(ext2[:, 3:], ext) and torch.equal(emb2, emb[:, 2:])
    # per-slot mask: bit 0 selects Out, bit 1 selects OutExtend (pull_box_extended_sparse_op.h:50-72)
    b2 = SlotBatch(keys, torch.tensor([0, 1, 2, 3, 4], dtype=torch.int32), 2, 2)      # 2 slots x 2 instances
    outs, outs_ext = pull_box_extended_sparse(b2, size=5, extend_size=2, box=box, mask=[0x02, 0x03])
    assert len(outs) == 1 and len(outs_ext) == 2
    assert torch.equal(outs[0], emb[2:4]) and torch.equal(outs_ext[0], ext[0:2]) and torch.equal(outs_ext[1], ext[2:4])
    with pytest.raises(ValueError):
        pull_box_extended_sparse(b2, size=5, extend_size=2, box=box, mask=[1])
    box.end_pass()
    # rank_attention2 = rank_attention
    gen = torch.Generator().manual_seed(1)
    xr = torch.rand(6, 5, generator=gen)
    ro = CO.build_rank_offset(torch.tensor([1, 1, 1, 2, 2, 3]), torch.tensor([1, 2, 3, 1, 2, 1]), 3)
    P = torch.rand(9 * 5, 4, generator=gen)
    assert torch.equal(CO.rank_attention2(xr, ro, P, 3), CO.rank_attention(xr, ro, P, 3))
    # fused_seq_tensor against an index-by-index restatement of the reference kernels
    ins, bc, L, sn, d, an, off = 2, 3, 4, 3, 2, 1, 1
    xs = torch.rand(ins, bc * sn * L * d, generator=gen) + 0.1
    xs.view(ins, bc, sn, L, d)[0, 0, :, 2:] = 0                     # two padded time steps
    ad = torch.rand(ins, bc * an * d, generator=gen)
    din, mask, side, sess = CO.fused_seq_tensor(xs, ad, bc, L, sn, d, an, off)
    assert din.shape == (bc, ins, L, 4 * an * d) and side.shape == (bc, ins, L, (sn - an) * d)
    assert sess.shape == (bc, ins, L, an * d) and mask.shape == (bc, ins, L)
    xf, af = xs.reshape(-1), ad.reshape(-1)
    one_slot, one_seq, piece = L * d, sn * L * d, an * d
    for b in range(bc):
        for i in range(ins):
            for t in range(L):
                for idx in range(piece):
                    sl, osl, k = idx // d + off, idx // d, idx % d
                    v = xf[i * bc * one_seq + b * one_seq + sl * one_slot + t * d + k]
                    a = af[i * bc * piece + b * piece + osl * d + k]
                    row = din[b, i, t]
                    assert row[osl * d + k] == v and row[piece + osl * d + k] == a
                    assert torch.isclose(row[2 * piece + osl * d + k], v - a) and torch.isclose(row[3 * piece + osl * d + k], v * a)
                    assert sess[b, i, t, osl * d + k] == v
                for idx in range((sn - an) * d):
                    sl, k = idx // d + 0, idx % d                    # ad_slot_offset != 0: side slots start at 0
                    assert side[b, i, t, idx] == xf[i * bc * one_seq + b * one_seq + sl * one_slot + t * d + k]
    assert mask[0, 0].tolist() == [1, 1, 0, 0] and mask[:, 1].min() == 1 and mask[1:].min() == 1
    din0, _, side0, _ = CO.fused_seq_tensor(xs, ad, bc, L, sn, d, an, 0)
    assert torch.equal(side0[0, 0, 0], xs.view(ins, bc, sn, L, d)[0, 0, 1:, 0].reshape(-1))     # offset 0: side = slots after
    # dedup contract
    k = torch.tensor([9, 3, 9, 1, 3, 9])
    merged, restore, sorted_idx, offset, cnts = R.dedup_keys_and_fillidx(k)
    assert merged.tolist() == [1, 3, 9] and cnts.tolist() == [1, 2, 3] and offset.tolist() == [0, 1, 3]
    assert torch.equal(merged[restore], k)
    for u in range(3):
        occ = sorted_idx[offset[u]:offset[u] + cnts[u]]
        assert torch.all(k[occ] == merged[u])


def test_data_generator_monitor_optimizer_fleet_and_predict_dataset(tmp_path, box, capsys):
    from paddlebox_b200 import fleet
    from paddlebox_b200.data_generator import MultiSlotDataGenerator
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.optimizer import Adam, BoxPSOptimizer
    from paddlebox_b200.utils import monitor

    class Gen(MultiSlotDataGenerator):
        def generate_sample(self, line):
            def it():
                for i in range(5):
                    yield [("label", [float(i % 2)]), ("dense", [0.5, 0.25]), ("s0", [i + 1, i + 2]), ("s1", [7])]
            return it

    g = Gen()
    g.set_batch(2)
    g.run_from_memory()
    text = capsys.readouterr().out
    assert text.count("\n") == 5 and text.splitlines()[0] == "1 0.0 2 0.5 0.25 2 1 2 1 7"
    out = io.StringIO()
    g.run_from_iterable(["x"], out)
    assert out.getvalue() == text
    (tmp_path / "d.txt").write_text(text * 8)
    monitor.stat_reset()
    monitor.stat_add("STAT_x", 2); monitor.stat_add("STAT_x")
    assert monitor.stat_get("STAT_x") == 3 and monitor.stat_all() == {"STAT_x": 3}
    monitor.stat_reset("STAT_x")
    assert monitor.stat_get("STAT_x", -1) == -1
    # BoxPSOptimizer.build_trainer + one pass + serving-style predict_dataset
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var(slot_vars(2, 2))
    ds.set_batch_size(8)
    ds.set_filelist([str(tmp_path / "d.txt")])
    ds.load_into_memory()
    assert ds.get_memory_data_size() == 40
    box.begin_feed_pass("20240101"); box.feed_pass(ds.unique_keys()); box.end_feed_pass(); box.begin_pass()
    torch.manual_seed(0)
    model = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=2, hidden=(8,))
    tr = BoxPSOptimizer(Adam(1e-2), sync_steps=1).build_trainer(model, box)
    assert tr.cfg.dense_lr == 1e-2 and model._param_need_sync
    stats = tr.train_from_dataset(ds)
    assert stats["instances"] == 40
    from paddlebox_b200.serving import Predictor
    pred = Predictor.__new__(Predictor)                  # reuse the live box: only the dataset loop is under test
    pred.box, pred.model = box, model
    ds.prepare_train(1, 0, shuffle=False)
    p = pred.predict_dataset(ds)
    assert p.shape == (40,) and np.all((p > 0) & (p < 1))
    box.end_pass()
    fleet.init()
    fleet.init_worker(); fleet.init_server()
    assert fleet.is_worker() and not fleet.is_server()
    fleet.util.print_on_rank("hello", 0)
    assert "hello" in capsys.readouterr().out


def test_lr_map_sets_per_parameter_rates_of_the_async_dense_table():
    import time
    from paddlebox_b200.host import AsyncDenseTable
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    t = AsyncDenseTable(8, lr=0.1, threads=2)
    t.set_params(torch.zeros(8))
    t.set_lr_range(4, 8, 0.0)                               # frozen half
    t.push(torch.ones(8)); t.wait()
    w = t.pull()
    assert torch.all(w[:4] < 0) and torch.all(w[4:] == 0)
    BoxWrapper.reset_instance()
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    try:
        box.initialize_gpu_and_load_model({}, lr_map={"fc_param_0": 0.0, "other": 0.5})
        m = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=2, hidden=(8,))
        names = [n for n, _ in m.named_dense_params()]
        box.lr_map = {"fc_param": 0.25, names[0]: 0.0}
        got = box.get_lr_map()
        assert got["fc_param.w_0"] == 0.25 and got["fc_param.b_0"] == 0.25 and got[names[0]] == 0.0
        tr = CTRTrainer(m, box, TrainerConfig(async_mode=True, dense_lr=0.05))
        p0 = dict(m.named_dense_params())[names[0]]
        others_before = tr.store.flat.clone()
        tr.store.gflat.fill_(1.0)
        tr.sync.after_backward()
        tr.sync.finalize()
        o = tr.store.offsets[[id(p) for p in tr.store.params].index(id(p0))]
        assert torch.equal(tr.store.flat[o:o + p0.numel()], others_before[o:o + p0.numel()])      # lr 0: untouched
        mask = torch.ones_like(tr.store.flat, dtype=torch.bool)
        mask[o:o + p0.numel()] = False
        assert (tr.store.flat[mask] != others_before[mask]).float().mean() > 0.5
    finally:
        BoxWrapper.reset_instance()


def test_donefile_manifest_and_resume(tmp_path, box):
    from paddlebox_b200.utils.donefile import ModelManifest, resume
    man = ModelManifest(str(tmp_path / "models"))
    assert man.latest() == (None, [])
    keys = np.arange(1, 21, dtype=np.uint64)
    rows = box.host_table.fetch_rows(keys)
    rows[:, 0], rows[:, 2] = 5, 1.0
    box.host_table.store_rows(keys, rows)
    box.save_delta(str(tmp_path / "d_old"))
    man.record_delta("20240101", 9, str(tmp_path / "d_old"))             # older than the base: ignored on resume
    box.save_base(str(tmp_path / "batch"), str(tmp_path / "xbox"), "20240102")
    man.record_base("20240102", str(tmp_path / "xbox"), str(tmp_path / "batch"))
    for p, w in ((1, 2.0), (2, 3.0)):
        rows[:5 * p, 2] = w
        box.host_table.store_rows(keys[:5 * p], rows[:5 * p])
        box.save_delta(str(tmp_path / f"d{p}"))
        man.record_delta("20240102", p, str(tmp_path / f"d{p}"))
    want = box.host_table.pull(keys)
    base, deltas = man.latest()
    assert base["day"] == "20240102" and [d["pass_id"] for d in deltas] == [1, 2]
    with open(os.path.join(man.root, "xbox_patch_done.txt"), "a") as f:
        f.write('{"id": "torn')                                            # a crash mid-append
    assert [d["pass_id"] for d in man.latest()[1]] == [1, 2]
    BoxWrapper.reset_instance()
    b2 = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    b2.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
    got = resume(b2, man)
    assert got["base"].endswith("batch") and len(got["deltas"]) == 2 and got["rows"] == 20 + 5 + 10
    assert torch.allclose(b2.host_table.pull(keys), want, atol=1e-6)
    BoxWrapper.reset_instance()
    b3 = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    b3.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
    assert resume(b3, man, serving=True)["base"].endswith("xbox")
    assert torch.allclose(b3.host_table.pull(keys), want, atol=1e-6)
    # dense parameters travel with the manifest entries
    from paddlebox_b200.checkpoint import save_persistables
    from paddlebox_b200.models.ctr import CtrDnn
    torch.manual_seed(1)
    m1 = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(4,))
    save_persistables(m1, str(tmp_path / "dense3"))
    man.record_delta("20240102", 3, str(tmp_path / "d2"), str(tmp_path / "dense3"))
    torch.manual_seed(2)
    m2 = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(4,))
    got = resume(b3, man, serving=True, model=m2)
    assert got["dense"].endswith("dense3")
    for (n1, p1), (n2, p2) in zip(m1.named_dense_params(), m2.named_dense_params()):
        assert n1 == n2 and torch.equal(p1, p2)


def test_fused_seqpool_concat_two_groups():
    from paddlebox_b200.ops.ctr_ops import fused_seqpool_concat
    gen = torch.Generator().manual_seed(0)
    B, S = 4, 3
    lens = torch.randint(0, 3, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    n = int(lod[-1])
    x1 = torch.rand(n, 6, generator=gen, requires_grad=True)
    x2 = torch.rand(n, 5, generator=gen)
    out = fused_seqpool_concat([x1, x2], S, offsets=[1, 2], dims=[3, 0], lod=lod, batch_size=B)
    assert out.shape == (B, S * (3 + 3))
    ll = lod.tolist()
    for s in range(S):
        for b in range(B):
            g = s * B + b
            want = torch.cat([x1[ll[g]:ll[g + 1], 1:4].sum(0), x2[ll[g]:ll[g + 1], 2:5].sum(0)])
            assert torch.allclose(out[b, s * 6:(s + 1) * 6], want, atol=1e-6)
    out.sum().backward()
    assert torch.all(x1.grad[:, 1:4] == 1) and torch.all(x1.grad[:, 0] == 0) and torch.all(x1.grad[:, 4:] == 0)
    pooled = torch.arange(24.).view(2, 12)
    sel = fused_seqpool_concat(pooled, 3, offsets=[0, 2], dims=[1, 2])
    assert sel.shape == (2, 3 + 6) and sel[0].tolist() == [0, 4, 8, 2, 3, 6, 7, 10, 11]


def test_batch_fc_layouts():
    from paddlebox_b200.ops.ctr_ops import batch_fc
    gen = torch.Generator().manual_seed(0)
    bc, ins, fin, out = 3, 5, 4, 2
    x2 = torch.rand(ins, bc * fin, generator=gen)
    w2 = torch.rand(fin, bc * out, generator=gen)
    b2 = torch.rand(1, bc * out, generator=gen)
    y = batch_fc(x2, w2, b2, batchcount=bc)
    for k in range(bc):
        want = x2[:, k * fin:(k + 1) * fin] @ w2[:, k * out:(k + 1) * out] + b2[:, k * out:(k + 1) * out]
        assert torch.allclose(y[:, k * out:(k + 1) * out], want, atol=1e-6)
    x3 = torch.rand(bc, ins, fin, generator=gen)
    yt = batch_fc(x3, w2, b2, batchcount=bc, transpose_weight=True, act="relu")
    for k in range(bc):
        want = torch.relu(x3[k] @ w2[:, k * out:(k + 1) * out] + b2[:, k * out:(k + 1) * out])
        assert torch.allclose(yt[k], want, atol=1e-6)
    with pytest.raises(ValueError):
        batch_fc(x3, w2, b2, transpose_weight=True)


def test_partial_ops_ranges():
    from paddlebox_b200.ops.ctr_ops import fused_concat, partial_concat, partial_sum
    a, b = torch.arange(12.).reshape(3, 4), torch.arange(12., 24.).reshape(3, 4)
    assert torch.equal(partial_concat([a, b], -3, 2), torch.cat([a[:, 1:3], b[:, 1:3]], 1))
    assert torch.equal(partial_concat([a, b], 2), torch.cat([a[:, 2:], b[:, 2:]], 1))
    assert torch.equal(partial_sum([a, b], -1), a[:, 3:] + b[:, 3:])
    for bad in ((4, 1), (-5, 1), (2, 3)):
        with pytest.raises(ValueError):
            partial_concat([a, b], *bad)
    assert torch.equal(fused_concat([a, b], -2, 2), torch.cat([a[:, :2], b[:, :2]], 1))     # fused_concat clamps to 0


def test_dataset_edge_cases_do_not_crash(tmp_path):
    from paddlebox_b200 import _native as N
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var(slot_vars(2, 1))
    ds.set_batch_size(4)
    ds.set_thread(2)
    # no files at all, then an empty file: an empty pass, no batches, nothing to save
    ds.set_filelist([])
    ds.load_into_memory()
    assert ds.get_memory_data_size() == 0
    (tmp_path / "empty").write_text("")
    ds.set_filelist([str(tmp_path / "empty")])
    ds.load_into_memory()
    ds.prepare_train(2)
    assert ds.get_memory_data_size() == 0 and list(ds.batches(worker=0)) == [] and list(ds.batches(worker=1)) == []
    with pytest.raises(N.NativeError):
        ds.save_pass_file(str(tmp_path / "none.arena"))
    # a missing file is an error, not a crash; the dataset stays usable
    ds.set_filelist([str(tmp_path / "missing")])
    with pytest.raises(N.NativeError):
        ds.load_into_memory()
    # malformed lines are counted and skipped; fewer instances than workers gives zero batches
    (tmp_path / "mixed").write_text("1 1.0 1 0.5 1 7 1 8\nnot a line\n1 0 1 0.5 0 1 9\n1 1.0 1 0.25 2 3 4 1 5\n\n")
    ds.set_filelist([str(tmp_path / "mixed")])
    ds.load_into_memory()
    assert ds.get_memory_data_size() == 2 and "bad_lines=3" in ds.stats()
    ds.prepare_train(4)
    assert all(list(ds.batches(worker=w)) == [] for w in range(4))
    ds.prepare_train(1, 0, shuffle=False)
    (b,) = list(ds.batches(worker=0))
    assert b.batch_size == 2 and sorted(b.keys.tolist()) == [3, 4, 5, 7, 8]
    with pytest.raises(N.NativeError):
        ds._native.pack_batch(0, 5, 0, 0, 0, False)            # batch index out of range
    # a pass file that is not one
    (tmp_path / "junk.arena").write_bytes(b"x" * 300)
    with pytest.raises(N.NativeError, match="pass file"):
        ds.load_pass_file(str(tmp_path / "junk.arena"))
    assert ds.get_memory_data_size() == 0                      # the previous pass was released first


def test_host_table_and_metric_edge_cases(tmp_path):
    from paddlebox_b200 import _native as N
    from paddlebox_b200.host import HostAuc, HostTable, sort_unique_u64, partition_by_owner
    t = HostTable(4)
    empty = np.zeros(0, np.uint64)
    assert t.fetch_rows(empty).shape == (0, t.layout.stride) and t.pull(empty).shape == (0, 7)
    t.push(empty, np.zeros((0, t.layout.push_stride), np.float32))
    assert t.size() == 0 and t.save(str(tmp_path / "m")) == 0 and t.shrink() == 0 and t.keys().size == 0
    # key 0 is the padding key: pulls zeros, pushes are ignored
    t.fetch_rows(np.array([0, 5], np.uint64))
    p = np.zeros((2, t.layout.push_stride), np.float32)
    p[:, 1] = 1
    t.push(np.array([0, 5], np.uint64), p)
    rows = t.fetch_rows(np.array([0, 5], np.uint64), create=False)
    assert rows[0, 0] == 0 and rows[1, 0] == 1 and torch.all(t.pull(np.array([0], np.uint64)) == 0)
    # unknown keys pull zeros and are not created; a push to an unknown key is dropped
    assert torch.all(t.pull(np.array([77], np.uint64)) == 0) and t.size() == 2
    t.push(np.array([77], np.uint64), p[:1])
    assert t.size() == 2
    # malformed checkpoint lines are skipped
    d = tmp_path / "ck"
    d.mkdir()
    (d / "part-000-00000").write_text("garbage\n12 0 0 3 1 0.5 0 2 0\n13 1\n\n")
    t2 = HostTable(4)
    assert t2.load(str(d)) == 1 and t2.keys().tolist() == [12]
    with pytest.raises((N.NativeError, OSError)):
        t2.save("/proc/definitely/not/writable")
    assert sort_unique_u64(empty).size == 0 and [x.size for x in partition_by_owner(empty, 3)] == [0, 0, 0]
    # metrics with no data, one class only, NaN predictions
    a = HostAuc(1000)
    out = a.compute()
    assert len(out) == 8 and out[7] == 0 and all(np.isfinite(out))
    a.add(np.array([0.2, 0.8], np.float32), np.array([1.0, 1.0], np.float32))
    out = a.compute()
    assert out[7] == 2 and np.isfinite(out[0])
    a.add(np.array([np.nan, np.inf], np.float32), np.array([0.0, 1.0], np.float32))
    assert a.nan_inf() == (1, 1)


def test_numeric_gradient_checks_of_composed_ops():
    """The reference's OpTest harness checks analytic against numeric gradients; same here in float64."""
    from torch.autograd import gradcheck
    from paddlebox_b200.ops import ctr_ops as CO
    gen = torch.Generator().manual_seed(0)

    def rnd(*shape):
        return torch.rand(*shape, generator=gen, dtype=torch.float64, requires_grad=True)

    ro = CO.build_rank_offset(torch.tensor([1, 1, 1, 2, 2]), torch.tensor([1, 2, 3, 2, 1]), 3)
    assert gradcheck(lambda x, p: CO.rank_attention(x, ro, p, 3), (rnd(5, 4), rnd(9 * 4, 3)))
    assert gradcheck(lambda x, w, b: CO.batch_fc(x, w, b), (rnd(3, 4, 5), rnd(3, 5, 2), rnd(3, 2)))
    assert gradcheck(lambda x, w, b: CO.batch_fc(x, w, b, batchcount=3), (rnd(4, 6), rnd(2, 9), rnd(1, 9)))
    assert gradcheck(lambda a, b: CO.partial_sum([a, b], 1, 2), (rnd(3, 4), rnd(3, 4)))
    assert gradcheck(lambda a, b: CO.partial_concat([a, b], -2), (rnd(3, 4), rnd(3, 4)))
    lod = torch.tensor([0, 2, 2, 3, 5, 6, 6], dtype=torch.int32)         # S = 2, B = 3
    assert gradcheck(lambda a, b: CO.fused_seqpool_concat([a, b], 2, [1, 0], [2, 2], lod=lod, batch_size=3),
                     (rnd(6, 4), rnd(6, 3)))
    assert gradcheck(lambda x, ad: CO.fused_seq_tensor(x, ad, 2, 3, 3, 2, 1, 1)[0], (rnd(2, 2 * 3 * 3 * 2), rnd(2, 2 * 1 * 2)))
    assert gradcheck(lambda x: CO.filter_by_instag(x, [0, 1, 2, 3], torch.tensor([1, 2, 1]), [1])[0], (rnd(3, 4),))
    assert gradcheck(lambda a: CO.fused_concat([a, a * 2], 1, 2), (rnd(3, 4),))


def test_layers_namespace_has_the_reference_names():
    from paddlebox_b200 import layers
    names = ("_pull_box_sparse _pull_box_extended_sparse _pull_gpups_sparse _store_q_value lookup_input "
             "pull_cache_value data_norm continuous_value_model fused_seqpool_cvm fused_seqpool_cvm_with_diff_thres "
             "fused_seqpool_cvm_with_conv fused_seqpool_cvm_with_pcoc fused_seqpool_cvm_tradew "
             "fused_seqpool_cvm_with_credit cross_norm_layer_hadamard scaled_fc scaled_int8fc fused_seqpool_concat "
             "fused_concat fused_seq_tensor rank_attention batch_fc shuffle_batch partial_concat partial_sum "
             "tdm_child tdm_sampler filter_by_instag").split()
    assert [n for n in names if not callable(getattr(layers, n, None))] == []
    from paddlebox_b200.side_tables import GpuReplicaCache, InputTable
    c = GpuReplicaCache(2)
    c.add_items([1.0, 2.0]); c.add_items([3.0, 4.0])
    c.to_hbm(torch.device("cpu"))
    assert layers.pull_cache_value(c, torch.tensor([1, 0])).tolist() == [[3.0, 4.0], [1.0, 2.0]]
    t = InputTable(1)
    t.add_index_data("k", [9.0])
    assert layers.lookup_input(t, torch.tensor([1, 0])).tolist() == [[9.0], [0.0]]


def test_distribute_transpiler_shim_follows_test_boxps():
    import torch.nn as nn
    from paddlebox_b200.optimizer import (Adam, BoxPSOptimizer, DistributeTranspiler, DistributeTranspilerConfig,
                                          GradAllReduce, LocalSGD)
    prog = nn.Linear(2, 1)
    BoxPSOptimizer(Adam(1e-3)).minimize(prog)
    cfg = DistributeTranspilerConfig()
    cfg.collective_mode = "single_process_multi_thread"
    assert DistributeTranspiler(cfg).transpile(trainer_id=0, program=prog, trainers="127.0.0.1:6174") == "allreduce"
    cfg.collective_mode = "grad_allreduce"
    with pytest.raises(ValueError):
        DistributeTranspiler(cfg).transpile(trainer_id=0, program=prog, trainers="127.0.0.1:6174")
    assert DistributeTranspiler(cfg).transpile(0, program=prog, trainers="a:1,b:2") == "allreduce"
    cfg.collective_mode, cfg.local_sgd_steps = "local_sgd", 4
    assert DistributeTranspiler(cfg).transpile(0, program=prog, trainers="a:1,b:2") == "kstep"
    assert prog._pipeline_opt.sync_dense_mode == "kstep" and prog._pipeline_opt.sync_weight_step == 4
    cfg.collective_mode = "nope"
    with pytest.raises(ValueError):
        DistributeTranspiler(cfg).transpile(0, program=prog)
    assert GradAllReduce(0).transpile(main_program=prog, rank=1, endpoints="x", current_endpoint="x") == "allreduce"
    assert LocalSGD(0).transpile(main_program=prog) == "kstep"


def test_cpu_binding_of_reader_and_trainer_threads(tmp_path, box, monkeypatch):
    """FLAGS_enable_binding_train_cpu: reader threads on the read-ins cores, the trainer on its share."""
    import threading
    from paddlebox_b200.flags import FLAGS
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.parallel.cluster import get_readins_cores, get_train_cores
    from paddlebox_b200.trainer import CTRTrainer
    assert FLAGS.enable_binding_train_cpu
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var(slot_vars(2, 1))
    ds.set_thread(3)
    ds.set_filelist(write_files(str(tmp_path), n_files=3, lines_per_file=50, n_sparse=2, n_dense=1, key_space=20))
    ds.load_into_memory()                                    # pinned readers still read everything
    assert ds.get_memory_data_size() == 150
    if len(os.sched_getaffinity(0)) < 4:
        pytest.skip("needs >= 4 cores to split")
    before = os.sched_getaffinity(threading.get_native_id())
    try:
        assert CTRTrainer(CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(4,)), box).pinned_cores == []
        monkeypatch.setenv("LOCAL_WORLD_SIZE", "2")
        monkeypatch.setenv("LOCAL_RANK", "1")
        tr = CTRTrainer(CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(4,)), box)
        train = get_train_cores()
        assert tr.pinned_cores == train[1::2] and os.sched_getaffinity(threading.get_native_id()) == set(train[1::2])
        assert not set(tr.pinned_cores) & set(get_readins_cores()) or len(before) < 2
    finally:
        os.sched_setaffinity(threading.get_native_id(), before)


def test_capacity_planner_for_the_headline_config():
    from paddlebox_b200.utils.capacity import plan
    # BASELINE config 3: 125 M keys / GPU, dim 16, adagrad, ~320 K keys per batch, 8 GPUs
    b = plan(125_000_000, 16, "adagrad", max_keys_per_batch=1 << 19, world_size=8, dense_params=600_000)
    assert b.fits() and 15e9 < b.table_bytes < 40e9 and b.engine_bytes < 2e9
    assert "GB HBM" in b.report()
    # the same table at dim 64 with per-element Adam state does not fit a 180 GB device any more at 450 M keys
    big = plan(450_000_000, 64, "adam", max_keys_per_batch=1 << 22, world_size=8)
    assert not big.fits() and big.hbm_bytes > 180e9
    one = plan(1000, 8, world_size=1, max_keys_per_batch=4096)
    assert one.mailbox_cap == 4096 and one.table_capacity == 1334


def test_lod_tensor_bytes_are_the_reference_layout():
    """Byte-exact check of the serialized LoDTensor (lod_tensor.cc:205-231, tensor_util.cc:1056-1110), hand
    assembled: u32 0 | u64 lod levels | u32 0 | i32 desc size | TensorDesc{data_type=FP32(5), dims} | data."""
    import struct
    from paddlebox_b200.checkpoint import deserialize_lod_tensor, serialize_lod_tensor
    t = torch.tensor([[1.0, 2.0]])
    desc = bytes([0x08, 0x05, 0x10, 0x01, 0x10, 0x02])
    want = struct.pack("<IQ", 0, 0) + struct.pack("<Ii", 0, len(desc)) + desc + struct.pack("<2f", 1.0, 2.0)
    assert serialize_lod_tensor(t) == want
    back = deserialize_lod_tensor(want)
    got = back[0] if isinstance(back, tuple) else back
    assert torch.equal(torch.as_tensor(got).reshape(1, 2), t)
    # one LoD level [0, 2, 3]: u64 byte size (24) then the three u64 offsets
    t3 = torch.arange(3, dtype=torch.int64)
    raw = serialize_lod_tensor(t3, lod=[[0, 2, 3]])
    head = struct.pack("<IQ", 0, 1) + struct.pack("<Q3Q", 24, 0, 2, 3)
    desc3 = bytes([0x08, 0x03, 0x10, 0x03])
    assert raw == head + struct.pack("<Ii", 0, len(desc3)) + desc3 + struct.pack("<3q", 0, 1, 2)


def test_data_norm_slot_dim_and_scale_shift():
    from paddlebox_b200.ops.functional import DataNorm
    torch.manual_seed(0)
    x = torch.rand(6, 8) + 0.5
    x[1, 0] = 0.0                                   # slot block 0 of row 1 has show == 0
    x[4, 4] = 0.0                                   # slot block 1 of row 4
    dn = DataNorm(8, slot_dim=4, enable_scale_and_shift=True)
    with torch.no_grad():
        dn.batch_sum.fill_(5e3)                      # mean 0.5, scale 1
        dn.scale_w.copy_(torch.arange(1., 9.))
        dn.bias.fill_(0.25)
    xr = x.clone().requires_grad_(True)
    y = dn(xr)
    want = ((x - 0.5) * torch.arange(1., 9.) + 0.25)
    want[1, :4] = 0
    want[4, 4:] = 0
    assert torch.allclose(y, want, atol=1e-6)
    size0, sum0, sq0 = dn.batch_size.clone(), dn.batch_sum.clone(), dn.batch_square_sum.clone()
    y.sum().backward()
    g = torch.arange(1., 9.).repeat(6, 1)
    g[1, :4] = 0
    g[4, 4:] = 0
    assert torch.allclose(xr.grad, g)
    live = torch.ones(6, 8)
    live[1, :4] = 0
    live[4, 4:] = 0
    assert torch.allclose(dn.bias.grad, live.sum(0)) and torch.allclose(dn.scale_w.grad, ((x - 0.5) * live).sum(0), atol=1e-5)
    # summaries: +1 per column, mean / mean-square of the live rows only (+ count * eps)
    cnt = live.sum(0)
    assert torch.allclose(dn.batch_size, size0 * dn.decay + 1)
    assert torch.allclose(dn.batch_sum, sum0 * dn.decay + (x * live).sum(0) / cnt, rtol=1e-6)
    assert torch.allclose(dn.batch_square_sum, sq0 * dn.decay + (((x - 0.5) ** 2) * live).sum(0) / cnt + cnt * dn.epsilon, rtol=1e-6)
    # plain layer with scale and shift only
    d2 = DataNorm(8, enable_scale_and_shift=True)
    with torch.no_grad():
        d2.scale_w.fill_(2.0); d2.bias.fill_(1.0)
    assert torch.allclose(d2(x), x * 2 + 1, atol=1e-5)           # mean 0, scale 1 at initialisation
    with pytest.raises(ValueError):
        DataNorm(7, slot_dim=4)


def test_dump_main_program(tmp_path, box, monkeypatch):
    from paddlebox_b200.flags import FLAGS
    from paddlebox_b200.models.ctr import DeepFM
    from paddlebox_b200.trainer import CTRTrainer
    monkeypatch.chdir(tmp_path)
    FLAGS.enable_dump_main_program = True
    try:
        tr = CTRTrainer(DeepFM(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(4,)), box)
    finally:
        FLAGS.enable_dump_main_program = False
    text = (tmp_path / "device_0_ops.txt").read_text()
    assert "[modules]" in text and "DeepFM" in text and "fc_0.w_0" in text and "dense sync" in text
    assert tr.dump_main_program(str(tmp_path / "again.txt")).endswith("again.txt")


def test_debug_flags_are_wired(tmp_path, box, capsys):
    from paddlebox_b200.flags import FLAGS
    from paddlebox_b200.side_tables import GpuReplicaCache
    from paddlebox_b200.utils.dump import FieldDumper
    box.init_metric("AucCalculator", "m", "label", "pred", metric_phase=1, bucket_size=100)
    box.add_metric_data({"pred": torch.tensor([0.2, 0.9]), "label": torch.tensor([0., 1.])})
    FLAGS.enable_debug_print_metrics_info = True
    FLAGS.padbox_dump_debug_lineid = "ins_b"
    FLAGS.gpu_replica_cache_dim = 3
    try:
        box.get_metric_msg("m")
        assert "[metric m] auc=1" in capsys.readouterr().out
        d = FieldDumper(str(tmp_path / "dump"), ["pred"])
        d.dump_batch(["ins_a", "ins_b"], {"pred": torch.tensor([0.1, 0.7])})
        d.close()
        text = "".join(open(os.path.join(tmp_path / "dump", f)).read() for f in os.listdir(tmp_path / "dump"))
        assert "ins_b" in text and "ins_a" not in text
        assert GpuReplicaCache().emb_dim == 3
    finally:
        FLAGS.enable_debug_print_metrics_info = False
        FLAGS.padbox_dump_debug_lineid = ""
        FLAGS.gpu_replica_cache_dim = 8


def test_remaining_contrib_layer_names():
    import torch
    from paddlebox_b200 import layers as L
    torch.manual_seed(0)
    w = torch.randn(10, 4, requires_grad=True)
    ids = torch.tensor([1, 2, 9, 3, 9, 0])
    lod = torch.tensor([0, 2, 2, 6])
    out = L.fused_embedding_seq_pool(ids, lod, w, padding_idx=-1)          # -1 = id 9 is padding
    want = torch.stack([w[1] + w[2], torch.zeros(4), w[3] + w[0]])
    assert torch.allclose(out, want)
    out.sum().backward()
    assert torch.allclose(w.grad[9], torch.zeros(4)) and torch.allclose(w.grad[1], torch.ones(4))
    x, y = torch.randn(3, 4), torch.randn(3, 4)
    o, mid = L.fused_elemwise_activation(x, y, ["elementwise_add", "relu"])
    assert torch.allclose(o, x + torch.relu(y)) and torch.allclose(mid, torch.relu(y))
    o = L.fused_elemwise_activation(x, y, "scale,elementwise_mul", scale=0.5, save_intermediate_out=False)
    assert torch.allclose(o, 0.5 * x * y)
    yb = torch.randn(3)
    o = L.fused_elemwise_activation(x, yb, ["tanh", "elementwise_add"], axis=0, save_intermediate_out=False)
    assert torch.allclose(o, torch.tanh(x + yb[:, None]))
    lr = L.pow2_decay_with_linear_warmup(10, 110, 0.1, 0.01)
    assert abs(lr(5) - 0.05) < 1e-12 and abs(lr(10) - 0.1) < 1e-12 and abs(lr(60) - (0.09 * 0.25 + 0.01)) < 1e-12
    assert lr(110) == 0.01 and lr(500) == 0.01 and L.pow2_decay_with_linear_warmup(10, 110, 0.1, 0.01, step=10) == lr(10)
    # sparse_embedding: per-id embedx rows of the wrapper
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch
    BoxWrapper.reset_instance()
    try:
        box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0, "mf_initial_range": 0.1})
        keys = torch.tensor([11, 12, 0, 11], dtype=torch.int64)
        batch = SlotBatch(keys, torch.tensor([0, 2, 4], dtype=torch.int32), 2, 1, torch.zeros(2, 1))
        box.begin_feed_pass("20240101"); box.feed_pass(keys); box.end_feed_pass(); box.begin_pass()
        e = L.sparse_embedding(batch, [1 << 40, 4], padding_idx=12)
        assert e.shape == (4, 4) and torch.equal(e[0], e[3]) and torch.equal(e[1], torch.zeros(4))
        import pytest
        with pytest.raises(ValueError):
            L.sparse_embedding(batch, 8)
        box.end_pass()
    finally:
        BoxWrapper.reset_instance()


def test_dataset_surface_matches_both_reference_flavours():
    """Every public (and fleet-flavour ``_set_*``) method name of the reference's dataset classes exists here,
    and the added ones do what they say."""
    import re
    import paddlebox_b200.dataset as D
    from tests.util_data import slot_vars
    refs = ["/root/reference/python/paddle/fluid/dataset.py",
            "/root/reference/python/paddle/distributed/fleet/dataset/dataset.py"]
    import os
    if all(os.path.exists(r) for r in refs):
        skip = {"_check_use_var_with_data_generator", "_desc", "_generate_local_tables_unlock"}   # proto / pslib internals
        for path in refs:
            cur = None
            for ln in open(path):
                m = re.match(r"class (\w+)", ln)
                if m:
                    cur = m.group(1)
                m = re.match(r"    def (\w+)", ln)
                if m and cur and not m.group(1).startswith("__") and m.group(1) not in skip:
                    # (every concrete dataset here derives from InMemoryDataset, which carries the setters the
                    # reference keeps on DatasetBase)
                    cls = D.InMemoryDataset if cur == "DatasetBase" else getattr(D, cur)
                    assert hasattr(cls, m.group(1)), f"{cur}.{m.group(1)} ({path})"
    ds = D.DatasetFactory().create_dataset("InMemoryDataset")
    ds.init(batch_size=4, thread_num=3, use_var=slot_vars(2, 1), pipe_command="cat", input_type=1)
    ds.update_settings(batch_size=8, thread_num=2, merge_size=2, parse_ins_id=True, fs_name="hdfs://x", fs_ugi="u,p",
                       data_feed_type="SlotRecordInMemoryDataFeed", queue_num=4, fea_eval=True, candidate_size=77)
    assert (ds.batch_size, ds.thread_num, ds.fs_name, ds.data_feed_type, ds.parse_ins_id) == \
        (8, 2, "hdfs://x", "SlotRecordInMemoryDataFeed", True)
    ds._set_batch_size(16)
    assert ds.batch_size == 16
    import pytest
    with pytest.raises(TypeError):
        ds.update_settings(no_such_setting=1)
    ds.set_filelist(["a"])
    ds.set_thread(8)
    ds._prepare_to_run()
    assert ds.thread_num == 1                                      # no more readers than files
    ds._finish_to_run(); ds._dynamic_adjust_before_train(2); ds._dynamic_adjust_after_train()
    ds._set_trainer_num(4); ds._set_heter_ps(True); ds.set_graph_config({"walk_len": 8})
    assert ds.trainer_num == 4 and ds.graph_config["walk_len"] == 8

    class PS:
        date = None

        def set_date(self, d):
            self.date = d
    ps = PS()
    ds._set_use_ps_gpu(ps)
    ds.set_date("20240107")
    assert ps.date == "20240107"
    it = D.DatasetFactory().create_dataset("InputTableDataset")
    it.set_index_parser("libidx.so"); it.set_feed_type("InputTableDataFeed")
    assert it.index_parser == "libidx.so"


def test_uid_slot_gives_the_user_id_of_every_instance(tmp_path):
    import torch
    import paddlebox_b200.dataset as D
    from tests.util_data import slot_vars, write_files
    files = write_files(str(tmp_path), n_files=1, lines_per_file=12, n_sparse=3, n_dense=1, key_space=100)
    ds = D.DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var(slot_vars(3, 1)); ds.set_batch_size(4); ds.set_thread(1); ds.set_merge_thread_num(1)
    ds.set_filelist(files)
    ds._set_uid_slot("slot1")
    ds.load_into_memory()
    ds.prepare_train(1, shuffle=False)
    uids = torch.cat([ds.uid_of_batch(b) for b in ds.batches(device=torch.device("cpu"))])
    # label(1) + dense(1) = tokens 0..3; then slot0: n v..; slot1: n v...
    want = []
    for line in open(files[0]):
        t = line.split()
        p = 4
        n0 = int(t[p]); p += 1 + n0
        want.append(int(t[p + 1]))
    assert uids.tolist() == want


def test_data_feed_desc_text_round_trips():
    import paddlebox_b200.dataset as D
    from paddlebox_b200.dataset import dense_slot, sparse_slot
    a = D.DatasetFactory().create_dataset("PadBoxSlotDataset")
    a.set_use_var([dense_slot("label", 1), sparse_slot("s1")], [dense_slot("label", 1), dense_slot("d", 3), sparse_slot("s0"), sparse_slot("s1")])
    a.set_batch_size(48); a.set_thread(6); a.set_pipe_command('awk "{print}"'); a.set_rank_offset("rank_offset"); a.set_ads_offset("ads")
    a.set_pv_batch_size(12); a.set_input_type(1); a.set_sample_rate(0.5); a.set_uid_slot("s1"); a.set_feed_type("SlotPaddleBoxDataFeed")
    a.set_graph_config({"walk_len": 8, "meta_path": "u2i-i2u", "gpu_graph_training": True})
    text = a.desc()
    assert 'name: "SlotPaddleBoxDataFeed"' in text and "is_used: false" in text and 'uid_slot: "s1"' in text
    b = D.DatasetFactory().create_dataset("PadBoxSlotDataset")
    b.set_data_feed_desc(text)
    assert b.desc() == text
    assert (b.batch_size, b.thread_num, b.pipe_command, b.rank_offset_name, b.pv_batch_size, b.input_type, b.sample_rate) == \
        (48, 6, 'awk "{print}"', "rank_offset", 12, 1, 0.5)
    assert [s.name for s in b.all_slots] == ["label", "d", "s0", "s1"] and [s.is_used for s in b.all_slots] == [True, False, False, True]
    assert b.graph_config == {"walk_len": 8, "meta_path": "u2i-i2u", "gpu_graph_training": True}
    # the one-line form older scripts print still parses
    c = D.DatasetFactory().create_dataset("InMemoryDataset")
    c.set_data_feed_desc('batch_size: 7 thread_num: 2 pipe_command: "cat" multi_slot_desc { slots { name: "x" type: "float" '
                         'is_dense: true is_used: true shape: 4 } slots { name: "k" type: "uint64" is_used: true } }')
    assert c.batch_size == 7 and c.dense_dim == 4 and c.num_sparse_slots == 1
