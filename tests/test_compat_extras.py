"""AUC runner (slot shuffle feature evaluation), cluster/thread helpers, HeterPS-flavoured names,
masked_data_norm, store_q_value, merge_multi_models — all on CPU."""
import numpy as np
import pytest
import torch

from paddlebox_b200.auc_runner import AucRunner
from paddlebox_b200.box_wrapper import BoxWrapper
from paddlebox_b200.dataset import DatasetFactory
from paddlebox_b200.models.ctr import CtrDnn
from paddlebox_b200.ops import functional as F
from paddlebox_b200.parallel import cluster
from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
from tests.util_data import slot_vars, write_files


@pytest.fixture
def box():
    BoxWrapper.reset_instance()
    b = BoxWrapper.set_instance(embedx_dim=8, device=torch.device("cpu"), rank=0, world_size=1)
    b.initialize_gpu_and_load_model({"mf_create_thresholds": 1.0, "learning_rate": 0.1, "mf_learning_rate": 0.1})
    yield b
    BoxWrapper.reset_instance()


def _dataset(tmp_path, n_sparse=6, n_dense=3, lines=300):
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var(slot_vars(n_sparse, n_dense))
    ds.set_batch_size(32)
    ds.set_thread(2)
    ds.set_filelist(write_files(str(tmp_path), n_files=2, lines_per_file=lines, n_sparse=n_sparse, n_dense=n_dense,
                                key_space=40))
    return ds


def test_slots_shuffle_back_restores_records(tmp_path, box):
    ds = _dataset(tmp_path)
    ds.load_into_memory()
    ds.begin_pass()
    ds.prepare_train(1)
    before = [b.keys.clone() for b in ds.batches()]
    ds.slots_shuffle(["slot0", "slot3"])
    during = [b.keys.clone() for b in ds.batches()]
    ds.slots_shuffle_back(["slot0", "slot3"])
    after = [b.keys.clone() for b in ds.batches()]
    assert any(a.numel() != d.numel() or not torch.equal(a, d) for a, d in zip(before, during))
    assert all(torch.equal(a, b) for a, b in zip(before, after))
    ds.end_pass(False)


def test_auc_runner_ranks_the_informative_slot(tmp_path, box):
    torch.manual_seed(0)
    ds = _dataset(tmp_path)
    box.init_metric("AucCalculator", "auc", "label", "pred", metric_phase=1, bucket_size=100000)
    model = CtrDnn(num_slots=6, embedx_dim=8, dense_dim=3, hidden=(32, 16))
    tr = CTRTrainer(model, box, TrainerConfig(dense_lr=1e-2))
    for _ in range(8):
        ds.load_into_memory()
        ds.begin_pass()
        tr.train_from_dataset(ds)
        ds.end_pass(False)
        ds.release_memory()
    box.get_metric_msg("auc")
    ds.load_into_memory()
    ds.begin_pass()
    shows_before = box.host_table.fetch_rows(box.host_table.keys(), create=False)[:, 0].sum()
    runner = AucRunner(box, [["slot0"], ["slot4"]], metric="auc",
                       slot_list=[f"slot{i}" for i in range(6)])
    res = runner.run(tr, ds)
    ds.end_pass(False)
    assert res["base"] > 0.7
    drop0, drop4 = res["groups"][0]["drop"], res["groups"][1]["drop"]
    assert drop0 > 0.1 and drop0 > drop4 + 0.05          # the clicks depend on slot0 only
    # evaluation ran in test mode: nothing was pushed
    shows_after = box.host_table.fetch_rows(box.host_table.keys(), create=False)[:, 0].sum()
    assert shows_before == shows_after
    with pytest.raises(ValueError):
        AucRunner(box, [["nope"]], slot_list=["slot0"])


def test_cluster_single_process_and_thread_helpers():
    c = cluster.Cluster.ins()
    assert c.size() == 1 and c.rank() == 0 and c.nccl_rank(3, 8) == 3
    assert c.allreduce(5, cluster.OP_MAX) == 5
    assert np.allclose(c.allreduce_sum([1.0, 2.0]), [1.0, 2.0])
    c.barrier()
    hits = [0] * 16
    cluster.execute_func(16, lambda i: hits.__setitem__(i, i * i))
    assert hits == [i * i for i in range(16)]
    acc = np.zeros(1000)

    def fill(b, e):
        acc[b:e] += 1
    cluster.exec_range_func(1000, fill, parts=7)
    assert (acc == 1).all()
    tc, rc = cluster.get_train_cores(), cluster.get_readins_cores()
    assert tc and rc
    assert cluster.set_cpu_affinity(tc + rc) in (True, False)


def test_masked_data_norm_matches_definition():
    torch.manual_seed(0)
    x = torch.randn(10, 4, requires_grad=True)
    mask = torch.tensor([1, 0, 1, 1, 0, 1, 1, 1, 0, 1], dtype=torch.bool)
    bsize, bsum, bsq = torch.full((4,), 100.0), torch.randn(4) * 10, torch.full((4,), 400.0)
    got = {}
    y = F.masked_data_norm(x, mask, bsize, bsum, bsq, 1e-4, lambda a, b, c: got.update(size=a, sum=b, sq=c))
    mean, scale = bsum / bsize, torch.sqrt(bsize / bsq)
    exp = (x.detach() - mean) * scale
    exp[~mask] = 0
    assert torch.allclose(y, exp, atol=1e-6)
    y.sum().backward()
    gexp = scale.expand(10, 4).clone()
    gexp[~mask] = 0
    assert torch.allclose(x.grad, gexp)
    xs = x.detach()[mask]
    assert torch.allclose(got["size"], torch.ones(4))
    assert torch.allclose(got["sum"], xs.mean(0), atol=1e-6)
    assert torch.allclose(got["sq"], ((xs - mean) ** 2).mean(0) + 1e-4, atol=1e-5)
    # no selected row: all three summaries are zero
    F.masked_data_norm(x, torch.zeros(10, dtype=torch.bool), bsize, bsum, bsq, 1e-4,
                       lambda a, b, c: got.update(size=a)).sum().backward()
    assert float(got["size"].sum()) == 0.0


def test_store_q_value_and_gpups_names(box):
    from paddlebox_b200 import ps_gpu
    q = [torch.rand(4, 1), torch.rand(4, 1)]

    class B:
        pass
    b = B()
    host = F.store_q_value(q, batch=b, device_id=0)
    assert len(host) == 2 and torch.equal(b.q_values[1], q[1]) and F.stored_q_values(0) is host
    assert ps_gpu.PSGPUWrapper.get_instance() is box          # one PS per process
    assert issubclass(ps_gpu.PSGPUWorker, CTRTrainer) and callable(ps_gpu.pull_gpups_sparse)


def test_merge_multi_models(tmp_path, box):
    keys = np.arange(1, 21, dtype=np.uint64)
    rows = box.host_table.fetch_rows(keys, create=True).copy()
    rows[:, 0], rows[:, 1], rows[:, 2] = 10.0, 2.0, 0.5
    box.host_table.store_rows(keys, rows)
    box.host_table.save(str(tmp_path / "m1"))
    rows[:, 0], rows[:, 1], rows[:, 2] = 4.0, 1.0, -0.25
    box.host_table.store_rows(keys[:10], rows[:10])
    box.host_table.save(str(tmp_path / "m2"))
    box.host_table.clear()
    box.merge_multi_models([str(tmp_path / "m1"), str(tmp_path / "m2")], update_type=0)
    r = box.host_table.fetch_rows(keys, create=False)
    assert np.allclose(r[:10, 0], 4.0) and np.allclose(r[:10, 2], -0.25) and np.allclose(r[10:, 0], 10.0)
    box.host_table.clear()
    box.merge_multi_models([str(tmp_path / "m1"), str(tmp_path / "m2")], update_type=1)
    r = box.host_table.fetch_rows(keys, create=False)
    assert np.allclose(r[:10, 0], 14.0) and np.allclose(r[:10, 1], 3.0) and np.allclose(r[:10, 2], -0.25)


def test_psgpu_sparse_sgd_setters_reach_the_table(box):
    from paddlebox_b200 import ps_gpu
    w = ps_gpu.PSGPUWrapper.get_instance()
    assert isinstance(w, BoxWrapper)
    # the instance was created through BoxWrapper; the setters are plain functions over it
    ps_gpu.PSGPUWrapper.set_sparse_sgd(w, learning_rate=0.2, initial_g2sum=1.0)
    ps_gpu.PSGPUWrapper.set_embedx_sgd(w, mf_create_thresholds=2.0, mf_learning_rate=0.3)
    cfg = w.host_table.config.as_dict()
    assert abs(cfg["learning_rate"] - 0.2) < 1e-7 and abs(cfg["mf_learning_rate"] - 0.3) < 1e-7
    assert abs(cfg["mf_create_thresholds"] - 2.0) < 1e-7 and abs(cfg["clk_coeff"] - 1.0) < 1e-7
    # a push now uses the new learning rate: embed_w moves by lr * g / show * sqrt(g0/(g0+g2sum))
    keys = np.array([77], dtype=np.uint64)
    before = w.host_table.fetch_rows(keys, create=True).copy()
    push = torch.zeros(1, w.host_table.layout.push_stride)
    push[0, 1], push[0, 3] = 1.0, 1.0           # show 1, embed_g 1
    w.host_table.push(keys, push)
    after = w.host_table.fetch_rows(keys, create=False)
    assert abs((after[0, 2] - before[0, 2]) - 0.2) < 1e-5


def test_quant_feature_type_serving_models_and_pull(tmp_path):
    """feature_type 3: serving saves hold int16 embedx steps, serving loads dequantise, the batch model
    keeps fp32, and pull_box_sparse returns embedx on the int16 grid."""
    import os
    import numpy as np
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch, pull_box_sparse
    BoxWrapper.reset_instance()
    scale = 1.0 / 256
    box = BoxWrapper.set_instance(embedx_dim=4, feature_type=3, pull_embedx_scale=scale, device=torch.device("cpu"),
                                  rank=0, world_size=1)
    try:
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
        assert box.seqpool_quant_ratio == 256
        keys = np.arange(1, 41, dtype=np.uint64)
        rows = box.host_table.fetch_rows(keys)
        L = box.host_table.layout
        rs = np.random.RandomState(0)
        rows[:, 3:7] = rs.uniform(-1, 1, (40, 4)).astype(np.float32)
        rows[:, L.off_mf] = 4
        rows[:, 0] = 20
        box.host_table.store_rows(keys, rows, mark_dirty=True)
        box.save_base(str(tmp_path / "batch"), str(tmp_path / "xbox"))
        line = open(os.path.join(tmp_path / "xbox", sorted(os.listdir(tmp_path / "xbox"))[0])).readline().split()
        body = [f for name in sorted(os.listdir(tmp_path / "xbox")) for f in open(tmp_path / "xbox" / name)]
        assert len(body) == 40
        for ln in body:
            tail = ln.split()[-4:]
            assert all(t.lstrip("-").isdigit() and -32768 <= int(t) <= 32767 for t in tail), ln
        from paddlebox_b200.host import HostTable
        t2 = HostTable(4, "adagrad", box.config)
        t2.set_serving_quant(scale)
        assert t2.load(str(tmp_path / "xbox"), serving=True) == 40
        got = t2.fetch_rows(keys, create=False)[:, 3:7]
        assert np.abs(got - rows[:, 3:7]).max() <= scale / 2 + 1e-7
        assert np.allclose(got / scale, np.round(got / scale), atol=1e-4)
        t3 = HostTable(4, "adagrad", box.config)                       # the batch model is full precision
        assert t3.load(str(tmp_path / "batch")) == 40
        assert np.allclose(t3.fetch_rows(keys, create=False)[:, 3:7], rows[:, 3:7], rtol=1e-6)
        # pull records on the grid
        box.begin_feed_pass("20240101")
        box.feed_pass(keys)
        box.end_feed_pass()
        box.begin_pass()
        k = torch.from_numpy(keys.astype(np.int64))
        lod = torch.arange(41, dtype=torch.int32)
        emb = pull_box_sparse(SlotBatch(k, lod, 40, 1), box=box)
        ex = emb[:, 3:].detach().numpy()
        assert np.abs(ex - rows[:, 3:7]).max() <= scale / 2 + 1e-7 and np.allclose(ex / scale, np.round(ex / scale), atol=1e-4)
        box.end_pass()
    finally:
        BoxWrapper.reset_instance()


def test_multi_dim_box_variable_embedding_widths(tmp_path):
    """Slots of widths 4 / 8 / 4 / 8 / 16: split, per-width pull, columns in slot order, training,
    checkpoint round trip."""
    import numpy as np
    import torch
    from paddlebox_b200.models.ctr import MultiDimDnn
    from paddlebox_b200.multi_dim import MultiDimBox
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch
    from paddlebox_b200.ops.seqpool import SeqpoolAttrs
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    dims = [4, 8, 4, 8, 16]
    mb = MultiDimBox(dims, device=torch.device("cpu"), rank=0, world_size=1)
    mb.initialize({"mf_create_thresholds": 0.0, "mf_initial_range": 0.05})
    gen = torch.Generator().manual_seed(0)
    B, S = 48, 5
    lens = torch.randint(0, 4, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    slot_of = torch.repeat_interleave(torch.arange(S).repeat_interleave(B), lens)
    keys = torch.randint(1, 60, (int(lod[-1]),), generator=gen, dtype=torch.int64) + 1000 * slot_of
    dense = torch.zeros(B, 1)
    first = keys[lod[:B].long().clamp(max=keys.numel() - 1)]
    dense[:, 0] = (first % 2).float()
    batch = SlotBatch(keys, lod, B, S, dense)
    subs = mb.split(batch)
    assert sorted(subs) == [4, 8, 16] and sum(s.N for s in subs.values()) == keys.numel()
    for d, sub in subs.items():                       # every key went to the table of its slot's width
        assert all(dims[int(k) // 1000] == d for k in sub.keys.tolist())
        assert sub.lod.numel() == sub.num_slots * B + 1 and int(sub.lod[-1]) == sub.N
    mb.begin_feed_pass("20240101")
    mb.feed_pass_batch(batch)
    mb.end_feed_pass()
    mb.begin_pass()
    # recognisable rows: embed_w = slot index
    for d, box in mb.boxes.items():
        ks = box._pass_keys[box._pass_keys != 0]
        rows = box.host_table.fetch_rows(ks)
        rows[:, 2] = (ks // 1000).astype(np.float32)
        box.host_table.store_rows(ks, rows, mark_dirty=False)
    attrs = SeqpoolAttrs(use_cvm=False, cvm_offset=2)
    x = mb.embedding_pool(batch, torch.ones(B, 2), attrs)
    widths = [1 + d for d in dims]
    assert x.shape == (B, sum(widths))
    off = 0
    for s, w in enumerate(widths):                    # column block s holds slot s: pooled embed_w = s * bag length
        bag = lens.view(S, B)[s].float()
        assert torch.allclose(x[:, off].detach(), s * bag), s
        off += w
    torch.manual_seed(0)
    tr = CTRTrainer(MultiDimDnn(mb, hidden=(16, 8)), mb.boxes[4], TrainerConfig(dense_lr=2e-2))
    losses = [float(tr.train_step(batch).detach()) for _ in range(40)]
    assert np.mean(losses[-5:]) < 0.9 * np.mean(losses[:5]), (losses[:3], losses[-3:])
    mb.end_pass()
    mb.save_base(str(tmp_path / "batch"), str(tmp_path / "xbox"))
    mb2 = MultiDimBox(dims, device=torch.device("cpu"), rank=0, world_size=1)
    mb2.initialize({"mf_create_thresholds": 0.0})
    assert mb2.load_model(str(tmp_path / "batch")) == sum(b.host_table.size() for b in mb.boxes.values())
    for d in mb.boxes:
        ks = mb.boxes[d].host_table.keys()
        assert np.allclose(mb2.boxes[d].host_table.fetch_rows(ks, create=False),
                           mb.boxes[d].host_table.fetch_rows(ks, create=False), rtol=1e-5, atol=1e-7)
    mb.finalize()
    mb2.finalize()


def test_dataset_pybind_parity_methods_and_desc_roundtrip(tmp_path):
    import numpy as np
    from paddlebox_b200.dataset import DatasetFactory
    from tests.util_data import slot_vars, write_files
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var(slot_vars(3, 2))
    ds.set_batch_size(8)
    ds.set_thread_num(2)
    text = ds.desc()
    ds2 = DatasetFactory().create_dataset("InMemoryDataset")
    ds2.set_data_feed_desc(text)
    assert ds2.desc() == text and ds2.batch_size == 8 and ds2.thread_num == 2
    ds2.set_filelist(write_files(str(tmp_path), n_files=1, lines_per_file=40, n_sparse=3, n_dense=2, key_space=30))
    ds2.set_preload_thread_num(1)
    ds2.create_channel(); ds2.create_preload_readers()
    ds2.preload_into_memory(); ds2.destroy_preload_readers(); ds2.wait_preload_done()
    assert ds2.get_memory_data_size() == 40 and ds2.thread_num == 1
    ds2.set_generate_unique_feasigns(True, 4)
    tabs = ds2.generate_local_tables_unlock(0, 8, 1, 1)
    allk = np.concatenate(tabs)
    assert len(tabs) == 4 and len(np.unique(allk)) == allk.size == ds2.unique_keys().size
    assert all(np.all(t % np.uint64(4) == np.uint64(i)) for i, t in enumerate(tabs))
    ds2.set_trainer_num(2)
    ds2.create_readers()
    assert ds2._prepared_workers == 2
    ds2.destroy_readers()
    ds2.set_current_phase(1); ds2.preprocess_instance(); ds2.postprocess_instance()
    ds2.set_shuffle_by_uid(True); ds2.dynamic_adjust_channel_num(4); ds2.set_gpu_graph_mode(False)
    assert ds2.enable_pv_merge() is False and callable(ds2.merge_by_lineid)


def test_tdm_tree_and_sampling():
    import torch
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.ops.tdm import TDMTree
    items = [50, 10, 40, 20, 30, 60]                       # 6 leaves -> binary tree of height 4 (8 leaf slots)
    tree = TDMTree(items, branch=2)
    assert tree.height == 4 and tree.layer_size == [1, 2, 3, 6]
    path = tree.travel(torch.tensor([50, 20, 999]))
    assert path[0].tolist() == [1, 2, 4, 8] and path[1].tolist() == [1, 2, 5, 11] and path[2].tolist() == [0, 0, 0, 0]
    kids, leaf = tree.children(torch.tensor([1, 3, 6, 0]))
    assert kids.tolist() == [[2, 3], [6, 0], [12, 13], [0, 0]] and leaf.tolist() == [[False, False], [False, False], [True, True], [False, False]]
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    sampler = ds.tdm_sample("t", items, [1, 2, 3], start_sample_layer=1, seed=3)
    ids = torch.tensor([10, 60, 30, 999] * 50)
    nodes, labels, mask = sampler(ids)
    assert nodes.shape == labels.shape == mask.shape == (200, 3 + 6)
    assert labels[0].tolist() == [1, 0, 1, 0, 0, 1, 0, 0, 0]
    trav = tree.travel(ids)
    col = 0
    for l, neg in zip((1, 2, 3), (1, 2, 3)):
        lo, size = tree.layer_start[l] + 1, tree.layer_size[l]
        assert torch.equal(nodes[:, col], trav[:, l])
        negs, m = nodes[:, col + 1:col + 1 + neg], mask[:, col + 1:col + 1 + neg].bool()
        assert torch.all((negs[m] >= lo) & (negs[m] < lo + size))              # inside the layer
        assert not torch.any(negs == trav[:, l:l + 1]) or not torch.any((negs == trav[:, l:l + 1]) & m)
        col += 1 + neg
    assert not mask[3].any() and mask[0, 0] == 1                             # unknown item: all masked
    # layer 1 has 2 nodes: one negative available; layer 2 has 3 nodes: 2 negatives
    assert mask[0].tolist() == [1, 1, 1, 1, 1, 1, 1, 1, 1]


def test_box_wrapper_pybind_parity_names(tmp_path):
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    BoxWrapper.reset_instance()
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    try:
        box.initialize_gpu_and_load_model({})
        box.init_metric("ContinueMaskCalculator", "c", "label", "pred", mask_varname="mask", metric_phase=1)
        box.add_metric_data({"pred": torch.tensor([0.5, 2.0, 9.0]), "label": torch.tensor([1.0, 1.0, 0.0]),
                             "mask": torch.tensor([1, 1, 0])})
        mae, rmse, actual, predicted, size = box.get_continue_metric_msg("c")
        assert size == 2 and abs(mae - 0.75) < 1e-6 and abs(actual - 1.0) < 1e-6 and abs(predicted - 1.25) < 1e-6
        runner = box.initialize_auc_runner([["s1"], ["s2", "s3"]], thread_num=2, slot_list=["s1", "s2", "s3"])
        assert runner.eval_slot_set() == {"s1", "s2", "s3"} and box.auc_runner is runner
        box.set_dataset_name("train"); box.set_input_table_dim(3); box.release_pool()
        assert box.input_table.dim == 3 and "host table rows" in box.print_device_info()
        assert box.init_afs_api("", "", "") is True
        (tmp_path / "a.txt").write_text("12345")
        assert box.file_mgr.list_info(str(tmp_path)) == [(str(tmp_path / "a.txt"), 5)]
    finally:
        from paddlebox_b200.flags import FLAGS
        FLAGS.padbox_auc_runner_mode = False
        BoxWrapper.reset_instance()


def test_pipeline_by_section_matches_plain_training():
    """cut_list splits a Sequential into sections run as a threaded micro-batch pipeline; gradients equal
    those of the unsplit model."""
    import copy
    from collections import OrderedDict
    import torch
    import torch.nn as nn
    from paddlebox_b200.optimizer import Adam, BoxPSOptimizer
    torch.manual_seed(0)
    model = nn.Sequential(OrderedDict([("fc1", nn.Linear(6, 8)), ("act1", nn.ReLU()), ("fc2", nn.Linear(8, 4)),
                                       ("act2", nn.Tanh()), ("fc3", nn.Linear(4, 1))]))
    ref = copy.deepcopy(model)
    opt = BoxPSOptimizer(Adam(1e-3), cut_list=["act1", "act2"], place_list=["cpu", "cpu", "cpu"],
                         concurrency_list=[2, 1, 2], queue_size=2)
    opt.minimize(model)
    pipe = model._section_pipeline
    assert len(pipe.sections) == 3 and [len(s) for s in pipe.sections] == [2, 2, 1]
    x = torch.randn(32, 6)
    y = torch.randn(32, 1)
    mbs, tgts = list(x.split(4)), list(y.split(4))
    loss = pipe.forward_backward(mbs, nn.functional.mse_loss, tgts)
    ref_loss = nn.functional.mse_loss(ref(x), y)
    ref_loss.backward()
    assert abs(loss - float(ref_loss.detach())) < 1e-5
    for p, q in zip(pipe.parameters(), ref.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-6)


def test_examples_run_on_cpu(tmp_path):
    import subprocess
    import sys
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    r = subprocess.run([sys.executable, os.path.join(root, "examples", "train_ctr.py"), "--cpu", "--passes", "4",
                        "--lines", "512", "--out", str(tmp_path)], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("pass ")]
    assert len(lines) == 4 and "day end" in r.stdout
    aucs = [float(ln.split("auc ")[1].split()[0]) for ln in lines]
    assert aucs[-1] > aucs[0] and aucs[-1] > 0.58, aucs
    assert os.path.isdir(tmp_path / "batch_model") and os.path.isdir(tmp_path / "xbox_base")
    assert os.path.exists(tmp_path / "donefiles" / "xbox_base_done.txt") and os.path.isdir(tmp_path / "dense_base")
    r = subprocess.run([sys.executable, os.path.join(root, "examples", "serve.py"), "--cpu", "--job", str(tmp_path)],
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    preds = [float(v) for v in r.stdout.split("predictions:")[1].split()]
    assert len(preds) == 16 and all(0.0 < v < 1.0 for v in preds) and len(set(preds)) > 1
    r = subprocess.run([sys.executable, os.path.join(root, "examples", "graph_skipgram.py"), "--cpu", "--epochs", "3"],
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    ls = [float(ln.rsplit(" ", 1)[1]) for ln in r.stdout.splitlines() if ln.startswith("epoch")]
    assert len(ls) == 3 and ls[-1] < ls[0]


def test_online_training_example_trains_days_and_resumes(tmp_path):
    import subprocess
    import sys
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    out = str(tmp_path / "job")
    cmd = [sys.executable, os.path.join(root, "examples", "online_train.py"), "--cpu", "--out", out, "--lines", "256"]
    r = subprocess.run(cmd + ["--days", "20240101,20240102"], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    first = [float(x.split("auc=")[1]) for x in r.stdout.split("summary:")[1].splitlines()[0].split()]
    assert len(first) == 4 and first[-1] > first[0]
    assert len(open(os.path.join(out, "donefile.txt")).read().strip().splitlines()) == 2
    assert len(open(os.path.join(out, "xbox_patch_done.txt")).read().strip().splitlines()) == 4
    assert os.path.isdir(os.path.join(out, "20240102", "0", "dense")) and os.path.isdir(os.path.join(out, "20240101", "base"))
    # second run: finds the last batch model, skips the finished days, continues with the new one
    r = subprocess.run(cmd + ["--days", "20240101,20240102,20240103"], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "resume from" in r.stdout and "20240102/0/" in r.stdout
    second = [x for x in r.stdout.split("summary:")[1].splitlines()[0].split()]
    assert [x.split(":")[0] for x in second] == ["20240103/1", "20240103/2"]
    assert float(second[0].split("auc=")[1]) > first[0]            # it did not start from scratch
    assert len(open(os.path.join(out, "donefile.txt")).read().strip().splitlines()) == 3


def test_psgpu_wrapper_pybind_names(tmp_path):
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.ps_gpu import PSGPUWrapper
    BoxWrapper.reset_instance()
    ps = PSGPUWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    try:
        assert isinstance(ps, PSGPUWrapper)
        ps.init()
        ps.init_gpu_ps([0], {"mf_create_thresholds": 0.0})
        ps.set_slot_vector([1, 2, 3]); ps.set_slot_dim_vector([4, 4, 4]); ps.set_slot_offset_vector([0, 1, 2])
        ps.set_sparse_sgd(learning_rate=0.2); ps.set_embedx_sgd(mf_learning_rate=0.3)
        assert abs(ps.host_table.config.learning_rate - 0.2) < 1e-7 and abs(ps.host_table.config.mf_learning_rate - 0.3) < 1e-7
        d = str(tmp_path / "x")
        assert ps.mkdir(d) and ps.exist(d) and ps.touchz(d + "/a") and ps.list(d) == [d + "/a"]
        assert ps.mv(d + "/a", d + "/b") and ps.cat(d + "/b") == "" and ps.upload(d + "/b", d + "/c")
        assert ps.download(d + "/e", d + "/c") and ps.remove(d + "/b") and sorted(ps.list(d)) == [d + "/c", d + "/e"]
        ps.build_task(torch.arange(1, 20))
        assert ps.host_table.size() == 20
        ps.end_pass()
    finally:
        BoxWrapper.reset_instance()


def test_psgpu_pass_pipeline_overlaps_build_with_training(tmp_path):
    """data_ready -> buildcpu_ready -> gpu_free (ps_gpu_wrapper.cc:885-955): the CPU half of pass 2 is built while
    pass 1 is still "training"; begin_pass of pass 2 waits for end_pass of pass 1; errors of the build thread
    surface in begin_pass; PSGPUTrainer.run drives a dataset through the same pipeline."""
    import threading
    import time
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.ps_gpu import HeterContext, PSGPUTrainer, PSGPUWrapper
    from tests.util_data import slot_vars, write_files
    BoxWrapper.reset_instance()
    ps = PSGPUWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    try:
        ps.init_gpu_ps([0], {"mf_create_thresholds": 0.0})
        ps.start_build_thread()
        t1, t2 = HeterContext(keys=np.arange(1, 101, dtype=np.uint64)), HeterContext(keys=np.arange(50, 301, dtype=np.uint64))
        ps.submit(t1)
        ps.submit(t2)                            # queued while pass 1 has not even begun
        ps.begin_pass()
        assert ps.current_task is t1 and np.array_equal(ps._pass_keys, np.arange(0, 101, dtype=np.uint64))
        deadline = time.time() + 10
        while t2.rows is None and time.time() < deadline:  # pass 2's CPU build finishes during pass 1
            time.sleep(0.01)
        assert t2.rows is not None and t2.rows.shape == (252, ps.host_table.layout.stride)
        assert ps.host_table.size() == 301                  # keys 0..300 exist in the CPU table already
        started = []
        th = threading.Thread(target=lambda: (ps.begin_pass(), started.append(time.time())))
        th.start()
        time.sleep(0.3)
        assert not started                                  # the GPU is still owned by pass 1
        t_end = time.time()
        ps.end_pass()
        th.join(timeout=10)
        assert started and started[0] >= t_end and ps.current_task is t2
        assert {"pre_build_s", "build_pull_s", "build_gpu_s"} <= set(t2.timing)
        ps.end_pass()
        bad = HeterContext(keys="not keys")
        ps.submit(bad)
        with pytest.raises(Exception):
            ps.begin_pass()
        # a dataset through PSGPUTrainer.run
        ds = DatasetFactory().create_dataset("InMemoryDataset")
        ds.set_use_var(slot_vars(3, 2)); ds.set_batch_size(16); ds.set_thread(1)
        ds.set_filelist(write_files(str(tmp_path), n_files=1, lines_per_file=64, n_sparse=3, n_dense=2, key_space=30))
        torch.manual_seed(0)
        tr = PSGPUTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=2, hidden=(8,)), ps)
        stats = tr.run(ds)
        assert stats["instances"] == 64 and ps.current_task is None
        assert "fc_0.w_0" in tr.finalize()
    finally:
        BoxWrapper.reset_instance()
