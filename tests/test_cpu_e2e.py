"""BASELINE config 1: CTR-DNN, 26 sparse + 13 dense slots, emb dim 8, one process on CPU —
the full plumbing slice (dataset -> feed pass -> begin_pass -> train -> end_pass -> metrics ->
save/load) without a GPU."""
import os

import numpy as np
import pytest
import torch

from paddlebox_b200.box_wrapper import BoxWrapper, owner_of, mix64_np
from paddlebox_b200.dataset import DatasetFactory
from paddlebox_b200.models.ctr import CtrDnn, DeepFM, WideDeep, MMoE
from paddlebox_b200.trainer import CTRTrainer, TrainerConfig, Executor
from tests.util_data import slot_vars, write_files


@pytest.fixture
def box():
    BoxWrapper.reset_instance()
    b = BoxWrapper.set_instance(embedx_dim=8, device=torch.device("cpu"), rank=0, world_size=1)
    b.initialize_gpu_and_load_model({"mf_create_thresholds": 1.0, "learning_rate": 0.1, "mf_learning_rate": 0.1})
    yield b
    BoxWrapper.reset_instance()


def _dataset(tmp_path, n_sparse=26, n_dense=13, lines=300, files=2, bs=32, **kw):
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var(slot_vars(n_sparse, n_dense))
    ds.set_batch_size(bs)
    ds.set_thread(2)
    ds.set_filelist(write_files(str(tmp_path), n_files=files, lines_per_file=lines, n_sparse=n_sparse,
                                n_dense=n_dense, key_space=40, **kw))
    return ds


def test_ctr_dnn_cpu_end_to_end(tmp_path, box):
    torch.manual_seed(0)
    ds = _dataset(tmp_path)
    ds.set_date("20240101")
    box.init_metric("AucCalculator", "join_auc", "label", "pred", metric_phase=1, bucket_size=100000)
    model = CtrDnn(num_slots=26, embedx_dim=8, dense_dim=13, hidden=(64, 32))
    trainer = CTRTrainer(model, box, TrainerConfig(dense_lr=1e-2))
    aucs = []
    for p in range(9):
        ds.preload_into_memory()
        ds.wait_preload_done()
        assert ds.get_memory_data_size() == 600
        ds.begin_pass()
        stats = trainer.train_from_dataset(ds)
        ds.end_pass(False)
        assert stats["instances"] == 600
        aucs.append(box.get_metric_msg("join_auc"))
        ds.release_memory()
    assert aucs[0][7] == 600                       # size
    assert aucs[-1][0] > aucs[0][0] and aucs[-1][0] > 0.7     # the model learns the planted signal
    # the table learnt show/click statistics and created embedx for frequent keys
    keys = box.host_table.keys()
    rows = box.host_table.fetch_rows(keys, create=False)
    assert rows[:, 0].max() > 1 and (rows[:, box.host_table.layout.off_mf] > 0).any()


def test_save_base_delta_and_resume(tmp_path, box):
    torch.manual_seed(0)
    ds = _dataset(tmp_path / "data" if (tmp_path / "data").mkdir() is None else tmp_path, n_sparse=4, n_dense=2,
                  lines=100, files=1)
    model = CtrDnn(num_slots=4, embedx_dim=8, dense_dim=2, hidden=(16,))
    trainer = CTRTrainer(model, box)
    ds.load_into_memory()
    ds.begin_pass()
    trainer.train_from_dataset(ds)
    ds.end_pass()
    msg = box.save_base(str(tmp_path / "batch_model"), str(tmp_path / "xbox_base"), "20240101")
    assert "batch_model=" in msg and "xbox_base=" in msg
    assert box.save_delta(str(tmp_path / "delta0")) == "xbox_delta=0"   # base save reset the baseline
    ds.begin_pass()
    trainer.train_from_dataset(ds)
    ds.end_pass(True, str(tmp_path / "delta1"))
    assert any(os.path.getsize(tmp_path / "delta1" / f) > 0 for f in os.listdir(tmp_path / "delta1"))
    # resume into a fresh wrapper: identical pull values
    keys = box.host_table.keys()
    before = box.host_table.pull(keys)
    box.save_base(str(tmp_path / "batch_model2"))
    BoxWrapper.reset_instance()
    b2 = BoxWrapper.set_instance(embedx_dim=8, device=torch.device("cpu"), rank=0, world_size=1)
    b2.initialize_gpu_and_load_model({}, model_path=str(tmp_path / "batch_model2"))
    assert torch.allclose(b2.host_table.pull(keys), before, rtol=1e-6, atol=1e-7)
    kept = b2.shrink_table(decay=0.5, delete_threshold=0.3)
    assert 0 < kept <= len(keys)


@pytest.mark.parametrize("cls,kw", [(WideDeep, {}), (DeepFM, {}), (MMoE, dict(num_experts=3, expert_hidden=(16,),
                                                                              tower_hidden=(8,), num_tasks=1))])
def test_other_model_families_step_on_cpu(tmp_path, box, cls, kw):
    torch.manual_seed(1)
    ds = _dataset(tmp_path, n_sparse=5, n_dense=3, lines=64, files=1, bs=32)
    ds.load_into_memory()
    ds.begin_pass()
    model = cls(num_slots=5, embedx_dim=8, dense_dim=3, **({"hidden": (16, 8)} if cls is not MMoE else {}), **kw)
    trainer = CTRTrainer(model, box)
    losses = []
    for _ in range(6):
        ds.prepare_train(1, 0, shuffle=False)
        for batch in ds.batches():
            losses.append(float(trainer.train_step(batch).detach()))
    ds.end_pass()
    assert all(np.isfinite(losses)) and np.mean(losses[-2:]) < np.mean(losses[:2])


def test_test_mode_does_not_update_table(tmp_path, box):
    ds = _dataset(tmp_path, n_sparse=3, n_dense=1, lines=50, files=1, bs=25)
    ds.load_into_memory()
    ds.begin_pass()
    model = CtrDnn(num_slots=3, embedx_dim=8, dense_dim=1, hidden=(8,))
    ex = Executor()
    keys = box.host_table.keys()
    before = box.host_table.fetch_rows(keys, create=False).copy()
    ex.infer_from_dataset(model, ds)
    assert np.array_equal(box.host_table.fetch_rows(keys, create=False), before)
    ex.train_from_dataset(model, ds)
    assert not np.array_equal(box.host_table.fetch_rows(keys, create=False), before)


def test_owner_hash_matches_device_definition():
    keys = np.array([0, 1, 2, 12345678901234567, 2**63 + 5], dtype=np.uint64)

    def mix(k):
        m = (1 << 64) - 1
        k ^= k >> 33
        k = (k * 0xff51afd7ed558ccd) & m
        k ^= k >> 33
        k = (k * 0xc4ceb9fe1a85ec53) & m
        k ^= k >> 33
        return k
    assert mix64_np(keys).tolist() == [mix(int(k)) for k in keys]
    assert owner_of(keys, 8).tolist() == [(mix(int(k)) >> 40) % 8 for k in keys]


def test_serving_predictor_reproduces_training_predictions(tmp_path, box):
    """xbox base + delta + dense persistables -> Predictor gives the trainer's own predictions."""
    from paddlebox_b200.checkpoint import save_persistables
    from paddlebox_b200.serving import Predictor
    torch.manual_seed(0)
    (tmp_path / "data").mkdir()
    ds = _dataset(tmp_path / "data", n_sparse=4, n_dense=2, lines=128, files=1)
    model = CtrDnn(num_slots=4, embedx_dim=8, dense_dim=2, hidden=(16,))
    trainer = CTRTrainer(model, box)
    ds.load_into_memory()
    for p in range(2):
        ds.begin_pass()
        trainer.train_from_dataset(ds)
        ds.end_pass()
        if p == 0:
            box.save_base(str(tmp_path / "batch"), str(tmp_path / "xbox_base"), "20240101")
    box.save_delta(str(tmp_path / "xbox_delta"))
    save_persistables(model, str(tmp_path / "dense"))
    ds.begin_pass()
    ds.prepare_train(1, 0, shuffle=False)
    batches = list(ds.batches())
    want = [trainer.infer_step(b)[1]["pred"].clone() for b in batches]
    ds.end_pass()
    n_feat = box.host_table.size()
    BoxWrapper.reset_instance()
    torch.manual_seed(5)                                   # different init: the weights must come from disk
    pred = Predictor(CtrDnn(num_slots=4, embedx_dim=8, dense_dim=2, hidden=(16,)), str(tmp_path / "xbox_base"),
                     [str(tmp_path / "xbox_delta")], dense_dir=str(tmp_path / "dense"), embedx_dim=8,
                     device=torch.device("cpu"))
    try:
        assert pred.num_features == n_feat
        for b, w in zip(batches, want):
            assert torch.allclose(pred.predict(b), w, atol=1e-5)
        rows_before = pred.box.host_table.fetch_rows(pred.box.host_table.keys(), create=False).copy()
        unseen = batches[0]
        unseen.keys = unseen.keys + 10_000_000              # unknown feasigns: zeros, and nothing is created
        p = pred.predict(unseen)
        assert torch.isfinite(p).all() and pred.num_features == n_feat
        assert np.array_equal(pred.box.host_table.fetch_rows(pred.box.host_table.keys(), create=False), rows_before)
    finally:
        pred.close()


def test_next_pass_preloads_and_feeds_while_the_current_pass_trains(tmp_path, box):
    """The PaddleBox pipeline: two datasets alternate; pass N+1 is read, parsed and fed (keys routed)
    while pass N is still between begin_pass and end_pass."""
    torch.manual_seed(0)
    dss = []
    for i in range(2):
        d = tmp_path / f"p{i}"
        d.mkdir()
        ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
        ds.set_use_var(slot_vars(4, 2))
        ds.set_batch_size(32)
        ds.set_thread(2)
        dss.append((ds, d))
    model = CtrDnn(num_slots=4, embedx_dim=8, dense_dim=2, hidden=(16,))
    trainer = CTRTrainer(model, box, TrainerConfig(dense_lr=1e-2))
    box.init_metric("AucCalculator", "auc", "label", "pred", metric_phase=1, bucket_size=10000)

    def files_for(p, d):
        return write_files(str(d), n_files=1, lines_per_file=256, n_sparse=4, n_dense=2, key_space=60 + 20 * p, seed=p)

    cur, cur_dir = dss[0]
    cur.set_filelist(files_for(0, cur_dir))
    cur.set_date("20240101")
    cur.preload_into_memory()
    cur.wait_preload_done()
    seen = []
    for p in range(4):
        nxt, nxt_dir = dss[(p + 1) % 2]
        cur.begin_pass()
        in_hbm = set(box._pass_keys.tolist())
        nxt.set_filelist(files_for(p + 1, nxt_dir))
        nxt.preload_into_memory()                      # pass p+1 is read while pass p trains
        stats = trainer.train_from_dataset(cur)
        nxt.wait_preload_done()                        # ... and fed before pass p ends
        assert set(box._pass_keys.tolist()) != in_hbm  # the NEXT pass's keys are already known
        cur.end_pass(False)
        assert stats["instances"] == 256
        seen.append(box.get_metric_msg("auc")[0])
        cur.release_memory()
        cur = nxt
    # every key of all four passes was created in the table and trained
    assert box.host_table.size() > 4 * 60 and all(0.0 <= a <= 1.0 for a in seen)


def test_join_and_update_phases_share_one_table(tmp_path, box):
    """The two-program pattern of PaddleBox jobs: a join phase on PV-merged batches and an update phase on
    plain instances, each with its own model and its own metric, flipping the phase between them."""
    torch.manual_seed(0)
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var(slot_vars(4, 2))
    ds.set_batch_size(32)
    ds.set_thread(2)
    ds.set_parse_logkey(True)
    ds.set_enable_pv_merge(True)
    ds.set_filelist(write_files(str(tmp_path), n_files=1, lines_per_file=192, n_sparse=4, n_dense=2, key_space=40,
                                logkey=True))
    ds.set_date("20240101")
    box.init_metric("AucCalculator", "join_auc", "label", "pred", metric_phase=1, bucket_size=10000)
    box.init_metric("AucCalculator", "update_auc", "label", "pred", metric_phase=0, bucket_size=10000)
    join = CTRTrainer(CtrDnn(num_slots=4, embedx_dim=8, dense_dim=2, hidden=(16,)), box, TrainerConfig(dense_lr=1e-2))
    update = CTRTrainer(CtrDnn(num_slots=4, embedx_dim=8, dense_dim=2, hidden=(8,)), box, TrainerConfig(dense_lr=1e-2))
    ds.load_into_memory()
    ds.begin_pass()
    for rnd in range(2):
        box.set_phase(1)                                   # join: page views stay together
        ds.set_current_phase(1)
        ds.preprocess_instance()
        ds.prepare_train(1, 0)
        sids = [b.meta["search_id"] for b in ds.batches(worker=0)]
        for s in sids:                                      # instances of one page view are adjacent
            uniq, counts = torch.unique_consecutive(s, return_counts=True)
            assert uniq.numel() == torch.unique(s).numel()
        st = join.train_from_dataset(ds)
        assert st["instances"] == 192
        ds.postprocess_instance()
        box.flip_phase()                                    # update: plain shuffled instances
        ds.set_current_phase(0)
        ds.prepare_train(1, 0)
        st = update.train_from_dataset(ds)
        assert st["instances"] == 192 and box.phase() == 0
        j, u = box.get_metric_msg("join_auc"), box.get_metric_msg("update_auc")
        assert j[7] == 192 and u[7] == 192                  # each metric saw only its own phase
    ds.end_pass(False)
    # both phases pushed into the same table: every occurrence was shown twice per round, two rounds
    rows = box.host_table.fetch_rows(box.host_table.keys(), create=False)
    total_occurrences = ds._native.total_feasigns()
    assert abs(rows[:, 0].sum() - 4 * total_occurrences) < 1e-3 * total_occurrences
