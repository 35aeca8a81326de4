"""End-to-end training on a real GPU: dataset -> pass lifecycle -> DeepFM step -> metrics, and
GPU-vs-CPU parity of the sparse path through the public API."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from paddlebox_b200.box_wrapper import BoxWrapper
from paddlebox_b200.dataset import DatasetFactory
from paddlebox_b200.models.ctr import CtrDnn, DeepFM, MMoE, WideDeep
from paddlebox_b200.ops.pull_box_sparse import SlotBatch, pull_box_sparse, sparse_embedding_pool, split_by_slot
from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
from tests.util_data import slot_vars, write_files


@pytest.fixture
def gbox():
    BoxWrapper.reset_instance()
    b = BoxWrapper.set_instance(embedx_dim=8, rank=0, world_size=1)
    b.initialize_gpu_and_load_model({"mf_create_thresholds": 1.0, "learning_rate": 0.1, "mf_learning_rate": 0.1},
                                    table_capacity=1 << 16, max_keys_per_batch=1 << 15)
    yield b
    BoxWrapper.reset_instance()


def _ds(tmp_path, n_sparse, n_dense, lines, bs, files=2):
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var(slot_vars(n_sparse, n_dense))
    ds.set_batch_size(bs)
    ds.set_thread(2)
    ds.set_filelist(write_files(str(tmp_path), n_files=files, lines_per_file=lines, n_sparse=n_sparse,
                                n_dense=n_dense, key_space=40))
    return ds


def test_deepfm_gpu_passes_learn(tmp_path, gbox):
    torch.manual_seed(0)
    ds = _ds(tmp_path, 26, 13, 300, 32)
    gbox.init_metric("AucCalculator", "auc", "label", "pred", metric_phase=1, bucket_size=100000)
    model = DeepFM(num_slots=26, embedx_dim=8, dense_dim=13, hidden=(64, 32)).cuda()
    tr = CTRTrainer(model, gbox, TrainerConfig(dense_lr=1e-2))
    aucs = []
    for p in range(9):
        ds.preload_into_memory()
        ds.wait_preload_done()
        ds.begin_pass()
        st = tr.train_from_dataset(ds)
        ds.end_pass()
        assert st["instances"] == 600
        aucs.append(gbox.get_metric_msg("auc"))
        ds.release_memory()
    assert aucs[0][7] == 600
    assert aucs[-1][0] > 0.7 and aucs[-1][0] > aucs[0][0]
    # pass write-back reached the host tier
    rows = gbox.host_table.fetch_rows(gbox.host_table.keys(), create=False)
    assert rows[:, 0].max() > 1


def test_gpu_sparse_path_matches_cpu(tmp_path):
    """Same data, same initial state: one pull + push through the public ops on GPU and CPU."""
    results = {}
    gen = torch.Generator().manual_seed(3)
    B, S, D = 16, 4, 8
    lens = torch.randint(0, 4, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    keys = torch.randint(1, 60, (int(lod[-1]),), generator=gen, dtype=torch.int64)
    dense = torch.rand(B, 3, generator=gen)
    dense[:, 0] = (dense[:, 0] < 0.4).float()
    w = torch.randn(S * (3 + D), generator=gen)
    for devname in ("cpu", "cuda"):
        BoxWrapper.reset_instance()
        dev = torch.device(devname, 0) if devname == "cuda" else torch.device("cpu")
        box = BoxWrapper.set_instance(embedx_dim=D, device=dev, rank=0, world_size=1)
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.5, "mf_initial_range": 0.0},
                                          table_capacity=1 << 12, max_keys_per_batch=4096)
        box.begin_feed_pass("20240101")
        box.feed_pass(keys)
        box.end_feed_pass()
        box.begin_pass()
        batch = SlotBatch(keys.to(dev), lod.to(dev), B, S, dense.to(dev))
        outs = []
        for step in range(3):
            cvm = torch.stack([torch.ones(B), dense[:, 0]], 1).to(dev)
            pooled = sparse_embedding_pool(batch, cvm, box)
            loss = (pooled * w.to(dev)).sum() / B
            loss.backward()
            outs.append(pooled.detach().cpu())
        # API-parity path: explicit embeddings
        emb = pull_box_sparse(batch, box=box)
        parts = split_by_slot(emb, batch)
        assert len(parts) == S and sum(p.shape[0] for p in parts) >= keys.numel()
        (emb.sum() / B).backward()
        if devname == "cuda":
            box.engine.check_status()
        box.end_pass()
        k = np.sort(box.host_table.keys())
        results[devname] = (outs, box.host_table.fetch_rows(k, create=False), k)
    BoxWrapper.reset_instance()
    for a, b in zip(results["cpu"][0], results["cuda"][0]):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-5)
    assert np.array_equal(results["cpu"][2], results["cuda"][2])
    slot_col = 4 + D + 1
    a, b = results["cpu"][1].copy(), results["cuda"][1].copy()
    a[:, slot_col] = b[:, slot_col] = 0          # slot id of multi-slot keys: last writer wins
    assert np.allclose(a, b, rtol=2e-3, atol=1e-4)


@pytest.mark.parametrize("cls,kw", [(CtrDnn, dict(hidden=(32, 16))), (WideDeep, dict(hidden=(32, 16))),
                                    (MMoE, dict(num_experts=3, expert_hidden=(32,), tower_hidden=(16,), num_tasks=2))])
def test_model_families_train_on_gpu(tmp_path, gbox, cls, kw):
    torch.manual_seed(1)
    n_lab = 2 if cls is MMoE else 1
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    vars_ = slot_vars(5, 3)
    ds.set_use_var(vars_)
    ds.set_batch_size(32)
    ds.set_filelist(write_files(str(tmp_path), n_files=1, lines_per_file=96, n_sparse=5, n_dense=3, key_space=30))
    ds.load_into_memory()
    ds.begin_pass()
    # MMoE reads two label columns: label + first dense feature thresholded (synthetic second task)
    model = cls(num_slots=5, embedx_dim=8, dense_dim=3 if n_lab == 1 else 2, **kw).cuda()
    tr = CTRTrainer(model, gbox, TrainerConfig(dense_lr=5e-3))
    losses = []
    for _ in range(8):
        ds.prepare_train(1, 0, shuffle=False)
        for batch in ds.batches():
            if n_lab == 2:
                batch.dense[:, 1] = (batch.dense[:, 1] > 0.5).float()
            losses.append(float(tr.train_step(batch).detach()))
    gbox.engine.check_status()
    ds.end_pass()
    assert all(np.isfinite(losses)) and np.mean(losses[-3:]) < np.mean(losses[:3])


def test_graft_smoke():
    import __graft_entry__ as ge
    ge.smoke()


@pytest.mark.parametrize("graph", [False, True])
def test_pipelined_feed_matches_plain_loop(tmp_path, graph):
    """dataset.host_batches -> CTRTrainer.train_host_batches (copy stream + batch_build into the
    step's buffers + loss ring) gives the same per-step losses as the plain loader loop."""
    def run(pipelined):
        BoxWrapper.reset_instance()
        box = BoxWrapper.set_instance(embedx_dim=8, rank=0, world_size=1)
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 1.0}, table_capacity=1 << 16,
                                          max_keys_per_batch=1 << 14)
        torch.manual_seed(3)
        ds = _ds(tmp_path, 6, 4, 128, 32, files=2)     # 256 instances -> 8 full batches
        ds.set_thread(1)                               # one reader and one merger: the record order
        ds.set_merge_thread_num(1)                     # (hence every batch) is reproducible
        model = DeepFM(num_slots=6, embedx_dim=8, dense_dim=4, hidden=(32, 16)).cuda()
        tr = CTRTrainer(model, box, TrainerConfig(dense_lr=1e-2))
        ds.load_into_memory()
        ds.begin_pass()
        ds.prepare_train(1)
        if graph:
            first = next(iter(ds.batches()))
            assert tr.enable_cuda_graph(first, 1 << 14, warmup=1), getattr(tr, "_graph_error", None)
        if pipelined:
            st = tr.train_host_batches(ds.host_batches())
            losses = st["losses"]
            assert st["h2d_bytes"] > 0 and st["d2h_bytes"] == 4 * st["steps"]
        else:
            losses = [float(tr.train_step(b).detach()) for b in ds.batches()]
        box.engine.check_status()
        ds.end_pass()
        BoxWrapper.reset_instance()
        return losses
    a, b = run(False), run(True)
    assert len(a) == len(b) == 8
    assert np.allclose(a, b, rtol=2e-3, atol=2e-4), (a, b)
