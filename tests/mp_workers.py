"""Worker functions for the multi-process tests (spawned with torch.multiprocessing)."""
import os
import traceback

import numpy as np
import torch
import torch.distributed as dist


def _init(rank, world, port, backend):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["RANK"] = str(rank)
    os.environ["WORLD_SIZE"] = str(world)
    if backend == "nccl":
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)


def run(fn, rank, world, port, backend, q, *args):
    try:
        _init(rank, world, port, backend)
        out = fn(rank, world, *args)
        q.put((rank, "ok", out))
    except Exception:
        q.put((rank, "error", traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


# ---------------------------------------------------------------- CPU / gloo
def cpu_feed_pass_and_dense(rank, world):
    from paddlebox_b200.box_wrapper import BoxWrapper, owner_of
    from paddlebox_b200.metrics import create_metric
    from paddlebox_b200.parallel.dense import DenseOptimizer, DenseParamStore, DenseSync
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({})
    # every rank saw different keys; after the feed pass each rank holds exactly the keys it owns
    mine = np.arange(rank * 100 + 1, rank * 100 + 181, dtype=np.uint64)
    box.begin_feed_pass("20240101")
    box.feed_pass(mine)
    box.end_feed_pass()
    allk = np.unique(np.concatenate([np.arange(r * 100 + 1, r * 100 + 181, dtype=np.uint64) for r in range(world)]
                                    + [np.zeros(1, np.uint64)]))
    expect = allk[owner_of(allk, world) == rank]
    assert np.array_equal(box._pass_keys, expect), (len(box._pass_keys), len(expect))
    # dense data parallel over gloo: grads averaged, replicas stay identical
    torch.manual_seed(0)
    model = torch.nn.Linear(8, 3)
    store = DenseParamStore(model, rank, world)
    opt = DenseOptimizer(store, lr=0.1, kind="sgd")
    sync = DenseSync(store, opt, "allreduce")
    x = torch.full((4, 8), float(rank + 1))
    model(x).sum().backward()
    g_local = store.gflat.clone()
    sync.after_backward()
    gathered = [torch.zeros_like(store.flat) for _ in range(world)]
    dist.all_gather(gathered, store.flat)
    assert all(torch.equal(gathered[0], t) for t in gathered)
    # k-step averaging
    sync2 = DenseSync(store, opt, "kstep", sync_weight_step=2)
    for _ in range(2):
        model(torch.randn(4, 8) * (rank + 1)).sum().backward()
        sync2.after_backward()
    dist.all_gather(gathered, store.flat)
    assert all(torch.allclose(gathered[0], t) for t in gathered)
    # ZeRO-1 sharding keeps replicas identical too
    store.gflat.zero_()
    sync3 = DenseSync(store, DenseOptimizer(store, lr=0.01, kind="adam"), "sharding")
    model(torch.randn(4, 8)).sum().backward()
    sync3.after_backward()
    dist.all_gather(gathered, store.flat)
    assert all(torch.allclose(gathered[0], t) for t in gathered)
    # async dense mode: one host table per rank, replicas meet again at the end of the pass
    store.gflat.zero_()
    sync4 = DenseSync(store, DenseOptimizer(store, lr=0.01, kind="adam"), "async")
    for _ in range(3):
        sync4.before_forward()
        model(torch.randn(4, 8) * (rank + 1)).sum().backward()
        sync4.after_backward()
    sync4.finalize()
    dist.all_gather(gathered, store.flat)
    assert all(torch.allclose(gathered[0], t, atol=1e-6) for t in gathered)
    # metric reduction: global AUC over both ranks' data
    m = create_metric("AucCalculator", name="a", bucket_size=10000)
    rs = np.random.RandomState(rank)
    pred = torch.from_numpy(rs.rand(500).astype(np.float32))
    label = (torch.from_numpy(rs.rand(500).astype(np.float32)) < pred).float()
    m.add_data({"pred": pred, "label": label})
    msg = m.get_metric_msg(dist.group.WORLD)
    assert msg[7] == 500 * world
    return msg[0]


# ---------------------------------------------------------------- GPU / nccl + peer memory
def gpu_sparse_and_dense(rank, world):
    from paddlebox_b200.box_wrapper import BoxWrapper, owner_of
    from paddlebox_b200.ops.sparse_table import SparseTable
    from paddlebox_b200.parallel.symm import SymmetricBuffer
    from paddlebox_b200.ops import reference as R
    dev = torch.device("cuda", rank)
    box = BoxWrapper.set_instance(embedx_dim=8, rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0, "mf_initial_range": 0.0},
                                      table_capacity=1 << 14, max_keys_per_batch=1 << 13)
    gen = torch.Generator().manual_seed(100 + rank)
    B, S, H = 32, 4, 11
    lens = torch.randint(0, 5, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    keys = torch.randint(1, 800, (int(lod[-1]),), generator=gen, dtype=torch.int64)
    box.begin_feed_pass("20240101")
    box.feed_pass(keys)
    box.end_feed_pass()
    # give every owned key a recognisable value: embed_w = key % 97, show = 50
    owned = box._pass_keys
    rows = box.host_table.fetch_rows(owned, create=True)
    rows[:, 0] = 50
    rows[:, 2] = (owned % 97).astype(np.float32)
    box.host_table.store_rows(owned, rows, mark_dirty=False)
    box.begin_pass()
    dk, dl = keys.to(dev), lod.to(dev)
    vals, restore = box.pull_sparse(dk)
    torch.cuda.synchronize()
    box.engine.check_status()
    got = vals[restore.long()][:, :H].cpu()
    assert torch.equal(got[:, 2], (keys % 97).float()), "pulled embed_w does not match the owner's value"
    assert torch.all(got[:, 0] == 50)
    # push: show counts must arrive at the owners (sum over ranks)
    dout = torch.zeros(B, S * H, device=dev)
    cvm = torch.ones(B, 2, device=dev)
    box.engine.push_pooled(restore, dout, cvm, dl, B, S, H, 2, 0, -float(B))
    torch.cuda.synchronize()
    box.engine.check_status()
    dist.barrier()
    box.end_pass()
    # gather every rank's keys to compute the expected global show increments
    all_keys = [None] * world
    dist.all_gather_object(all_keys, keys.numpy())
    cnt = {}
    for ks in all_keys:
        for k in ks.tolist():
            cnt[k] = cnt.get(k, 0) + 1
    rows2 = box.host_table.fetch_rows(owned, create=False)
    for k, r in zip(owned.tolist(), rows2):
        if k == 0:
            continue
        assert abs(r[0] - (50 + cnt.get(k, 0))) < 1e-4, (k, r[0], cnt.get(k, 0))
    # symmetric-memory allreduce / reduce-scatter / all-gather / broadcast
    n = 100003
    buf = SymmetricBuffer(n, rank, world)
    data = torch.arange(n, device=dev, dtype=torch.float32) * (rank + 1)
    buf.tensor.copy_(data)
    buf.allreduce(0, n, 1.0 / world)
    torch.cuda.synchronize()
    expect = torch.arange(n, device=dev, dtype=torch.float32) * sum(range(1, world + 1)) / world
    assert torch.allclose(buf.tensor, expect, rtol=1e-6)
    buf.tensor.fill_(float(rank))
    buf.collective("broadcast", 0, n, root=world - 1)
    torch.cuda.synchronize()
    assert torch.all(buf.tensor == world - 1)
    buf.check_status()
    return True


def gpu_deepfm_data_parallel(rank, world):
    """Two data-parallel ranks: replicas stay bit-identical, loss decreases."""
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.models.ctr import DeepFM
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    dev = torch.device("cuda", rank)
    box = BoxWrapper.set_instance(embedx_dim=8, rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0}, table_capacity=1 << 14,
                                      max_keys_per_batch=1 << 13)
    gen = torch.Generator().manual_seed(7 + rank)
    B, S = 64, 6
    lens = torch.randint(1, 4, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    keys = torch.randint(1, 300, (int(lod[-1]),), generator=gen, dtype=torch.int64)
    dense = torch.rand(B, 4, generator=gen)
    dense[:, 0] = (keys[lod[:B].long()] % 2).float()       # label depends on the first key of slot 0
    box.begin_feed_pass("20240101")
    box.feed_pass(keys)
    box.end_feed_pass()
    box.begin_pass()
    torch.manual_seed(rank)    # different init on purpose: the trainer must broadcast rank 0's weights
    model = DeepFM(num_slots=S, embedx_dim=8, dense_dim=3, hidden=(32, 16)).to(dev)
    tr = CTRTrainer(model, box, TrainerConfig(dense_lr=1e-2))
    batch = SlotBatch(keys.to(dev), lod.to(dev), B, S, dense.to(dev))
    losses = [float(tr.train_step(batch).detach()) for _ in range(30)]
    torch.cuda.synchronize()
    box.engine.check_status()
    gathered = [torch.zeros_like(tr.store.flat) for _ in range(world)]
    dist.all_gather(gathered, tr.store.flat)
    assert all(torch.equal(gathered[0], t) for t in gathered), "dense replicas diverged"
    assert np.mean(losses[-5:]) < np.mean(losses[:5])
    # the same step replayed from a CUDA graph keeps working across ranks
    ok = tr.enable_cuda_graph(batch, 1 << 13)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag) == 0:
        tr._graph = None
    l2 = [float(tr.train_step(batch).detach()) for _ in range(5)]
    torch.cuda.synchronize()
    box.engine.check_status()
    dist.all_gather(gathered, tr.store.flat)
    assert all(torch.equal(gathered[0], t) for t in gathered)
    box.end_pass()
    return (losses[0], losses[-1], bool(int(flag)), l2[-1])


def cpu_global_shuffle_and_cluster(rank, world):
    """Each rank loads its own files, the records are re-partitioned over the ranks by
    search_id, and the host-side Cluster collectives agree."""
    import tempfile
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.parallel.cluster import Cluster, OP_MAX, OP_SUM
    from tests.util_data import slot_vars, write_files
    base = os.path.join(tempfile.gettempdir(), f"pbx_gs_{os.environ['MASTER_PORT']}")
    os.makedirs(os.path.join(base, f"r{rank}"), exist_ok=True)
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var(slot_vars(4, 2))
    ds.set_batch_size(16)
    ds.set_thread(1)
    ds.set_parse_logkey(True)
    files = write_files(os.path.join(base, f"r{rank}"), n_files=1, lines_per_file=100 + 40 * rank, n_sparse=4,
                        n_dense=2, key_space=50, seed=rank, logkey=True)
    ds.set_filelist(files)
    ds.load_into_memory()
    n_before = ds.get_memory_data_size()
    ds.set_shuffle_dir(os.path.join(base, "x"))
    ds.global_shuffle(mode="search_id")
    n_after = ds.get_memory_data_size()
    # second round over the network transport (no shared directory): nothing lost either
    ds._shuffle_dir = None
    ds.global_shuffle(mode="random")
    n_net = torch.tensor([ds.get_memory_data_size()])
    dist.all_reduce(n_net)
    assert int(n_net) == 240, int(n_net)
    c = Cluster.ins()
    total_before = c.allreduce(int(n_before), OP_SUM)
    total_after = c.allreduce(int(n_after), OP_SUM)
    biggest = c.allreduce(int(n_after), OP_MAX)
    sums = c.allreduce_sum([float(rank + 1), 1.0])
    ranks = c.allgather_object(rank)
    from paddlebox_b200 import fleet
    fleet.init()
    assert fleet.worker_index() == rank and fleet.worker_num() == world and fleet.is_first_worker() == (rank == 0)
    shard = fleet.util.get_file_shard([f"f{i}" for i in range(5)])
    assert shard == ([f"f{i}" for i in range(3)] if rank == 0 else ["f3", "f4"])
    assert fleet.util.all_reduce(rank + 1, "max") == world and fleet.util.all_gather(rank) == list(range(world))
    c.barrier()
    if rank == 0:
        import shutil
        shutil.rmtree(base, ignore_errors=True)
    return dict(before=n_before, after=n_after, total_before=total_before, total_after=total_after,
                biggest=biggest, sums=sums.tolist(), ranks=ranks)


def cpu_collective_ops(rank, world):
    """c_allreduce_xsum / c_mixallgather (modes 0, 1, 2) / c_broadcast on CPU tensors over gloo."""
    from paddlebox_b200.parallel import collective as Cc
    a = torch.full((5,), float(rank + 1))
    b = torch.arange(6, dtype=torch.float32).reshape(2, 3) * (rank + 1)
    Cc.c_allreduce_xsum([a, b], scale=0.5)
    tot = sum(range(1, world + 1))
    ok = torch.allclose(a, torch.full((5,), 0.5 * tot)) and \
        torch.allclose(b, torch.arange(6, dtype=torch.float32).reshape(2, 3) * tot * 0.5)
    g = torch.full((7,), float(rank + 1))
    Cc.c_mixallgather([g], mode=0)                     # mean
    ok = ok and torch.allclose(g, torch.full((7,), tot / world))
    g1 = torch.full((7,), float(rank + 1))
    Cc.c_mixallgather([g1], mode=1)                    # reduce-scatter + all-gather == mean
    ok = ok and torch.allclose(g1, torch.full((7,), tot / world))
    # mode 2: every rank contributes its own contiguous shard of the tensor
    n = 9
    chunk = (n + world - 1) // world
    full = torch.arange(n, dtype=torch.float32)
    mine = torch.zeros(n)
    lo, hi = min(n, rank * chunk), min(n, rank * chunk + chunk)
    mine[lo:hi] = full[lo:hi]
    Cc.c_mixallgather([mine], mode=2)
    ok = ok and torch.equal(mine, full)
    w = torch.full((3,), float(rank))
    Cc.c_broadcast([w], root=1)
    ok = ok and torch.equal(w, torch.full((3,), 1.0))
    Cc.barrier()
    return bool(ok)


def cpu_multi_node_dense(rank, world):
    """4 ranks as 2 nodes x 2 devices: the three-leg dense sync gives what a flat allreduce gives."""
    from paddlebox_b200.parallel.dense import DenseOptimizer, DenseParamStore, DenseSync
    from paddlebox_b200.parallel.topology import NodeTopology
    topo = NodeTopology(gpus_per_node=2)
    assert (topo.nnodes, topo.local_size) == (2, 2) and topo.rank == topo.node_rank * 2 + topo.local_rank
    # raw three-leg allreduce on an odd-sized buffer (the last shard is short)
    g = torch.Generator().manual_seed(100 + rank)
    x = torch.randn(37, generator=g)
    ref = x.clone()
    dist.all_reduce(ref)
    topo.hierarchical_allreduce(x, 0.25)
    assert torch.allclose(x, ref * 0.25, atol=1e-6), (x - ref * 0.25).abs().max()
    # SyncDense: sum and all-gather across the nodes for one local rank
    s = torch.full((3,), float(rank))
    out = topo.sync_dense(s.clone())
    assert torch.equal(out, torch.full((3,), float(topo.local_rank * 2 + 2)))       # r + (r + 2)
    ag = topo.sync_dense(s, allgather=True)
    assert torch.equal(ag, torch.tensor([float(topo.local_rank)] * 3 + [float(topo.local_rank + 2)] * 3))
    from paddlebox_b200.parallel.collective import c_mixallgather
    ts = [torch.full((5,), float(rank + 1)), torch.full((2, 3), float(rank))]
    c_mixallgather(ts, mode=1, topo=topo)
    assert torch.allclose(ts[0], torch.full((5,), 2.5)) and torch.allclose(ts[1], torch.full((2, 3), 1.5))
    # the trainer-level modes keep all four replicas identical and equal to the single-node result
    res = {}
    for mode in ("allreduce", "kstep", "sharding"):
        finals = []
        for use_topo in (True, False):
            torch.manual_seed(0)
            model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 1))
            store = DenseParamStore(model, rank, world, topo=topo if use_topo else None)
            opt = DenseOptimizer(store, lr=0.05, kind="adam")
            sync = DenseSync(store, opt, mode, sync_weight_step=2)
            gen = torch.Generator().manual_seed(7 + rank)
            for _ in range(4):
                model(torch.randn(8, 6, generator=gen)).pow(2).mean().backward()
                sync.after_backward()
            sync.finalize()
            gathered = [torch.zeros_like(store.flat) for _ in range(world)]
            dist.all_gather(gathered, store.flat)
            assert all(torch.allclose(gathered[0], t, atol=1e-6) for t in gathered), mode
            finals.append(store.flat.clone())
        assert torch.allclose(finals[0], finals[1], atol=1e-5), (mode, (finals[0] - finals[1]).abs().max())
        res[mode] = float(finals[0].sum())
    return res


def cpu_sharded_sparse_router(rank, world):
    """CPU world: the table is sharded over the ranks' host tables, pull / push travel over gloo
    all-to-all, several ranks pushing one key are merged on its owner, a DeepFM trains on top."""
    from paddlebox_b200.box_wrapper import BoxWrapper, owner_of
    from paddlebox_b200.models.ctr import DeepFM
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch
    from paddlebox_b200.parallel.sparse_dist import owner_of_t
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    box = BoxWrapper.set_instance(embedx_dim=8, device=torch.device("cpu"), rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0, "mf_initial_range": 0.0})
    assert box.router is not None
    gen = torch.Generator().manual_seed(100 + rank)
    B, S, H = 32, 4, 11
    lens = torch.randint(0, 5, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    keys = torch.randint(1, 800, (int(lod[-1]),), generator=gen, dtype=torch.int64)
    assert np.array_equal(owner_of_t(keys, world).numpy(), owner_of(keys.numpy().view(np.uint64), world))
    box.begin_feed_pass("20240101")
    box.feed_pass(keys)
    box.end_feed_pass()
    owned = box._pass_keys
    assert np.all(owner_of(owned, world) == rank)
    rows = box.host_table.fetch_rows(owned, create=True)
    rows[:, 0] = 50
    rows[:, 2] = (owned % 97).astype(np.float32)
    box.host_table.store_rows(owned, rows, mark_dirty=False)
    box.begin_pass()
    n_before = box.host_table.size()
    vals, restore = box.pull_sparse(keys)
    got = vals[restore.long()]
    assert torch.equal(got[:, 2], (keys % 97).float()) and torch.all(got[:, 0] == 50)
    assert box.host_table.size() == n_before, "a pull created rows on a rank that does not own them"
    grads = torch.zeros(keys.numel(), H)
    grads[:, 0] = 1.0                                        # one show per occurrence
    batch = SlotBatch(keys, lod, B, S)
    box.push_sparse_grad(keys, restore, grads, batch.key2slot(), B)
    dist.barrier()
    all_keys = [None] * world
    dist.all_gather_object(all_keys, keys.numpy())
    cnt = {}
    for ks in all_keys:
        for k in ks.tolist():
            cnt[k] = cnt.get(k, 0) + 1
    rows2 = box.host_table.fetch_rows(owned, create=False)
    for k, r in zip(owned.tolist(), rows2):
        if k:
            assert abs(r[0] - (50 + cnt.get(k, 0))) < 1e-4, (k, r[0], cnt.get(k, 0))
    # data-parallel DeepFM over the routed table
    dense = torch.rand(B, 4, generator=gen)
    keys2 = keys.clone()
    lens2 = torch.clamp(lens, min=1)
    lod2 = torch.zeros(S * B + 1, dtype=torch.int32)
    lod2[1:] = torch.cumsum(lens2, 0)
    keys2 = torch.randint(1, 300, (int(lod2[-1]),), generator=gen, dtype=torch.int64)
    dense[:, 0] = (keys2[lod2[:B].long()] % 2).float()
    box.begin_feed_pass("20240102")
    box.feed_pass(keys2)
    box.end_feed_pass()
    box.begin_pass()
    torch.manual_seed(rank)
    model = DeepFM(num_slots=S, embedx_dim=8, dense_dim=3, hidden=(16, 8))
    tr = CTRTrainer(model, box, TrainerConfig(dense_lr=1e-2))
    b2 = SlotBatch(keys2, lod2, B, S, dense)
    losses = [float(tr.train_step(b2).detach()) for _ in range(25)]
    gathered = [torch.zeros_like(tr.store.flat) for _ in range(world)]
    dist.all_gather(gathered, tr.store.flat)
    assert all(torch.equal(gathered[0], t) for t in gathered), "dense replicas diverged"
    assert np.mean(losses[-5:]) < np.mean(losses[:5]), losses
    box.end_pass()
    return (losses[0], losses[-1])


def cpu_reshard_on_load(rank, world, ckpt):
    """A model saved by one process loads into `world` ranks: every rank keeps exactly the keys it owns."""
    from paddlebox_b200.box_wrapper import BoxWrapper, owner_of
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({})
    n = box.load_model(ckpt)
    keys = box.host_table.keys()
    assert n == keys.size and np.all(owner_of(keys, world) == rank)
    rows = box.host_table.fetch_rows(keys, create=False)
    assert np.allclose(rows[:, 2], (keys % 97).astype(np.float32))
    tot = torch.tensor([n])
    dist.all_reduce(tot)
    # save again from `world` ranks into one directory: the union is the original model
    out = ckpt + f"_resaved_{world}"
    box.save_base(out)
    dist.barrier()
    return int(tot)


def cpu_one_rank_per_node(rank, world):
    """Every rank is its own node (gpus_per_node = 1): the intra-node legs are trivial, the inter-node leg
    carries the whole buffer."""
    from paddlebox_b200.parallel.dense import DenseOptimizer, DenseParamStore, DenseSync
    from paddlebox_b200.parallel.topology import NodeTopology
    topo = NodeTopology(gpus_per_node=1)
    assert (topo.nnodes, topo.local_size, topo.local_rank, topo.node_rank) == (world, 1, 0, rank)
    x = torch.arange(10.) * (rank + 1)
    topo.hierarchical_allreduce(x, 1.0 / world)
    assert torch.allclose(x, torch.arange(10.) * sum(range(1, world + 1)) / world)
    torch.manual_seed(0)
    model = torch.nn.Linear(4, 2)
    store = DenseParamStore(model, rank, world, topo=topo)
    for mode in ("allreduce", "sharding"):
        sync = DenseSync(store, DenseOptimizer(store, lr=0.1, kind="sgd"), mode)
        model(torch.full((3, 4), float(rank + 1))).sum().backward()
        sync.after_backward()
        gathered = [torch.zeros_like(store.flat) for _ in range(world)]
        dist.all_gather(gathered, store.flat)
        assert all(torch.allclose(gathered[0], t) for t in gathered), mode
    return True


def cpu_router_edge_cases(rank, world):
    """One rank with no keys at all, heavy duplicates, the padding key 0, and keys nobody registered."""
    from paddlebox_b200.box_wrapper import BoxWrapper
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
    mine = torch.tensor([7, 7, 7, 0, 9, 7], dtype=torch.int64) if rank == 1 else torch.zeros(0, dtype=torch.int64)
    box.begin_feed_pass("20240101")
    box.feed_pass(mine)
    box.end_feed_pass()
    box.begin_pass()
    keys = torch.cat([mine, torch.tensor([12345], dtype=torch.int64)]) if rank == 1 else mine   # 12345 was never fed
    vals, restore = box.pull_sparse(keys)
    assert restore.numel() == keys.numel() and vals.shape[1] == box.hidden
    if rank == 1:
        got = vals[restore.long()]
        assert torch.all(got[3] == 0) and torch.all(got[6] == 0)           # key 0 and the unknown key pull zeros
        assert torch.equal(got[0], got[1]) and torch.equal(got[0], got[5])
    grads = torch.zeros(keys.numel(), box.hidden)
    grads[:, 0] = 1.0
    box.push_sparse_grad(keys, restore, grads, torch.zeros(keys.numel(), dtype=torch.int32), 4)
    dist.barrier()
    box.end_pass()
    total = torch.tensor([box.host_table.size()])
    dist.all_reduce(total)
    shows = 0.0
    ks = box.host_table.keys()
    if ks.size:
        shows = float(box.host_table.fetch_rows(ks, create=False)[:, 0].sum())
    s = torch.tensor([shows])
    dist.all_reduce(s)
    # keys 0 (registered by both ranks' feed pass on its owner), 7, 9 exist; 12345 was not created by the push
    assert int(total) in (3, 4) and abs(float(s) - 5.0) < 1e-5, (int(total), float(s))
    return True


def cpu_train_from_dataset_sharded(rank, world, data_dir):
    """Every rank loads ITS files (set_rank_world) and must train ALL of them, in the same number of
    steps as its peers although the files differ in size."""
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    from tests.util_data import slot_vars
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({})
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var(slot_vars(3, 2))
    ds.set_batch_size(16)
    ds.set_thread(1)
    ds.set_filelist(sorted(os.path.join(data_dir, f) for f in os.listdir(data_dir)))
    ds.set_rank_world(rank, world)
    ds.set_date("20240101")
    ds.load_into_memory()
    loaded = ds.get_memory_data_size()
    ds.begin_pass()
    torch.manual_seed(0)
    trainer = CTRTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=2, hidden=(8,)), box, TrainerConfig())
    stats = trainer.train_from_dataset(ds)
    ds.end_pass(False)
    return {"loaded": int(loaded), "instances": int(stats["instances"]), "steps": int(stats["steps"])}


def cpu_legacy_global_shuffle(rank, world):
    """Record-based global shuffle: every record travels to hash(ins_id) % world in the Record wire format."""
    from paddlebox_b200 import legacy_feed as LF
    ds = LF.MultiSlotDataset(["uint64", "float"], batch_size=4)
    ds.set_trainer(rank, world)
    ds.records = [LF.Record(np.array([1000 * rank + i], np.uint64), np.zeros(1, np.uint16), np.array([0.5], np.float32),
                            np.ones(1, np.uint16), f"ins{rank}_{i}") for i in range(50 + 10 * rank)]
    before = len(ds.records)
    after = ds.global_shuffle(seed=3)
    ids = sorted(r.ins_id for r in ds.records)
    import hashlib
    ok_owner = all(int.from_bytes(hashlib.md5(i.encode()).digest()[:8], "little") % world == rank for i in ids)
    tot = torch.tensor([before, after], dtype=torch.int64)
    dist.all_reduce(tot)
    return {"before": before, "after": after, "total_before": int(tot[0]), "total_after": int(tot[1]), "owner_ok": ok_owner,
            "sum_signs": int(sum(int(r.u64_signs[0]) for r in ds.records))}


def cpu_dense_overlap(rank, world):
    """Bucketed allreduce started from autograd hooks must give the plain allreduce's parameters (exactly in
    fp32, closely with the bf16 transport), start buckets DURING backward, and tolerate unused parameters."""
    from paddlebox_b200.flags import FLAGS, set_flags
    from paddlebox_b200.parallel.dense import DenseOptimizer, DenseParamStore, DenseSync

    def make():
        torch.manual_seed(0)
        return torch.nn.Sequential(torch.nn.Linear(16, 64), torch.nn.ReLU(), torch.nn.Linear(64, 64), torch.nn.ReLU(),
                                   torch.nn.Linear(64, 32), torch.nn.ReLU(), torch.nn.Linear(32, 1))

    def train(overlap_mb, bf16, steps=4, skip_last=False, kind="adam"):
        set_flags({"pbx_dense_overlap_mb": overlap_mb, "pbx_dense_overlap_bf16": bf16})
        model = make()
        store = DenseParamStore(model, rank, world)
        sync = DenseSync(store, DenseOptimizer(store, lr=0.05, kind=kind), "allreduce")
        g = torch.Generator().manual_seed(100 + rank)
        started = []
        for _ in range(steps):
            x = torch.randn(32, 16, generator=g)
            sync.before_forward()
            h = model[:-1](x) if skip_last else model(x)         # skip_last: the last layer gets no gradient
            h.pow(2).mean().backward()
            if sync.overlap is not None:
                started.append(sync.overlap.launched_in_backward)
            sync.after_backward()
        n_buckets = len(sync.overlap.buckets) if sync.overlap is not None else 0
        if sync.overlap is not None:
            sync.overlap.remove()
        return store.flat.clone(), n_buckets, started

    try:
        ref, _, _ = train(0.0, False)
        # 0.004 MB = 4 KB buckets: the 4-layer tower splits into several
        got, nb, started = train(0.004, False)
        assert nb >= 3, nb
        assert torch.equal(ref, got), float((ref - got).abs().max())
        assert all(s >= nb - 1 for s in started), (started, nb)      # all but (at most) the last bucket under backward
        # bf16 transport: compared under SGD (Adam's normalisation turns a rounding of a tiny gradient into a full step)
        ref_sgd, _, _ = train(0.0, False, kind="sgd")
        got16, _, _ = train(0.004, True, kind="sgd")
        assert torch.allclose(ref_sgd, got16, atol=2e-3) and not torch.equal(ref_sgd, got16), \
            float((ref_sgd - got16).abs().max())
        ref_skip, _, _ = train(0.0, False, skip_last=True)
        got_skip, _, started = train(0.004, False, skip_last=True)
        assert torch.equal(ref_skip, got_skip)
        # replicas stay identical
        gathered = [torch.zeros_like(got) for _ in range(world)]
        dist.all_gather(gathered, got)
        assert all(torch.equal(gathered[0], t) for t in gathered)
    finally:
        set_flags({"pbx_dense_overlap_mb": 0.0, "pbx_dense_overlap_bf16": False})
    return {"buckets": nb, "started": started}


def cpu_psgpu_pipeline_two_ranks(rank, world):
    """PSGPUWrapper's pre-build thread routes keys between the ranks while the main thread keeps issuing
    collectives on the world group (what training the previous pass does): the two must not interleave."""
    from paddlebox_b200.box_wrapper import owner_of
    from paddlebox_b200.ps_gpu import PSGPUWrapper
    PSGPUWrapper.reset_instance()
    box = PSGPUWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=rank, world_size=world)
    box.init_gpu_ps(conf={"mf_create_thresholds": 0.0})
    all_keys = [np.arange(1, 400, dtype=np.uint64) * np.uint64(r + 3) for r in range(world)]
    sizes = []
    for p in range(3):
        box.start_build_thread()
        from paddlebox_b200.ps_gpu import HeterContext
        task = HeterContext(keys=all_keys[rank] + np.uint64(1000 * p))
        box.submit(task)
        for i in range(25):                                  # "training": collectives on the world group meanwhile
            t = torch.full((64,), float(rank + i))
            dist.all_reduce(t)
            assert float(t[0]) == sum(r + i for r in range(world))
        box.begin_pass()
        mine = box._pass_keys
        assert mine.size and np.all(owner_of(mine, world) == rank)
        want = np.unique(np.concatenate([k + np.uint64(1000 * p) for k in all_keys] + [np.zeros(1, np.uint64)]))
        want_mine = want[owner_of(want, world) == rank]
        assert np.array_equal(mine, want_mine)
        sizes.append(int(mine.size))
        box.end_pass()
    box.finalize()
    PSGPUWrapper.reset_instance()
    return sizes


def cpu_queue_dataset_streaming_two_ranks(rank, world, data_dir):
    """Streaming dataset on two ranks with unequal file shares: both take the same number of steps (the pass
    ends for everyone when the first rank runs dry) and the sharded sparse path does not hang."""
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    from tests.util_data import slot_vars
    files = sorted(os.path.join(data_dir, f) for f in os.listdir(data_dir) if f.startswith("part-"))
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
    ds = DatasetFactory().create_dataset("QueueDataset")
    ds.set_use_var(slot_vars(3, 2)); ds.set_batch_size(10); ds.set_thread(1)
    ds.set_filelist(files)
    ds.set_rank_world(rank, world)
    assert len(ds.file_groups()) == (3 if rank == 0 else 2)
    torch.manual_seed(0)
    tr = CTRTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=2, hidden=(8,)), box, TrainerConfig())
    box.begin_pass()
    st = tr.train_from_dataset(ds)
    box.end_pass()
    flat = tr.store.flat.clone()
    gathered = [torch.zeros_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    assert all(torch.allclose(gathered[0], t) for t in gathered)      # dense replicas stayed in step
    BoxWrapper.reset_instance()
    return {"steps": st["steps"], "instances": st["instances"]}


def cpu_user_sharded_files_and_global_metrics(rank, world, data_dir):
    """The job-script idiom: every rank takes its files with ``fleet.util.get_file_shard`` (no set_rank_world) and
    reads metrics through the wrapper.  Every rank must train ALL the records it loaded, and the metric must
    cover the records of all ranks."""
    from paddlebox_b200 import fleet
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    from tests.util_data import slot_vars
    fleet.init()
    files = sorted(os.path.join(data_dir, f) for f in os.listdir(data_dir) if f.startswith("part-"))
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
    box.init_metric("AucCalculator", "auc", "label", "pred", metric_phase=1, bucket_size=10000)
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var(slot_vars(3, 2)); ds.set_batch_size(10); ds.set_thread(1)
    ds.set_filelist(fleet.util.get_file_shard(files))
    ds.load_into_memory()
    local = ds.get_memory_data_size()
    torch.manual_seed(0)
    tr = CTRTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=2, hidden=(8,)), box, TrainerConfig())
    ds.begin_pass()
    st = tr.train_from_dataset(ds)
    ds.end_pass(False)
    size = box.get_metric_msg("auc")[7]
    BoxWrapper.reset_instance()
    return {"local": local, "trained": st["instances"], "metric_size": int(size), "steps": st["steps"]}


def cpu_preload_overlap_two_ranks(rank, world, data_dir):
    """The production pipeline on several ranks: pass N+1 is read, parsed and fed (its keys routed to their
    owners over the all-to-all) while pass N is between begin_pass and end_pass; ranks stay in step and every
    key of every pass ends up on exactly one owner."""
    from paddlebox_b200.box_wrapper import BoxWrapper, owner_of
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    from tests.util_data import slot_vars, write_files
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
    box.init_metric("AucCalculator", "auc", "label", "pred", metric_phase=1, bucket_size=10000)
    torch.manual_seed(0)
    tr = CTRTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=2, hidden=(8,)), box, TrainerConfig())
    dss = []
    for i in range(2):
        ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
        ds.set_use_var(slot_vars(3, 2)); ds.set_batch_size(16); ds.set_thread(2)
        ds.set_rank_world(rank, world)
        dss.append(ds)

    def files_for(p):
        d = os.path.join(data_dir, f"p{p}")
        if rank == 0:
            os.makedirs(d, exist_ok=True)
            write_files(d, n_files=2 * world, lines_per_file=48 + 16 * (p % 2), n_sparse=3, n_dense=2,
                        key_space=50 + 10 * p, seed=p)
        dist.barrier()
        return sorted(os.path.join(d, f) for f in os.listdir(d))

    cur = dss[0]
    cur.set_filelist(files_for(0))
    cur.set_date("20240101")
    cur.preload_into_memory(); cur.wait_preload_done()
    sizes, steps = [], []
    for p in range(3):
        nxt = dss[(p + 1) % 2]
        cur.begin_pass()
        nxt.set_filelist(files_for(p + 1))
        nxt.preload_into_memory()
        st = tr.train_from_dataset(cur)
        nxt.wait_preload_done()                          # collective: keys of pass p+1 to their owners
        cur.end_pass(False)
        sizes.append(int(box.get_metric_msg("auc")[7]))
        steps.append(st["steps"])
        cur.release_memory()
        cur = nxt
    keys = box.host_table.keys()
    assert np.all(owner_of(keys[keys != 0], world) == rank)
    tot = torch.tensor([keys.size]); dist.all_reduce(tot)
    flat = tr.store.flat.clone()
    g = [torch.zeros_like(flat) for _ in range(world)]
    dist.all_gather(g, flat)
    assert all(torch.allclose(g[0], t) for t in g)
    BoxWrapper.reset_instance()
    return {"sizes": sizes, "steps": steps, "keys_total": int(tot)}


def cpu_wuauc_two_ranks(rank, world):
    """Per-user AUC over two ranks: users live on one rank each here, so the reduced result must equal the
    single-process result on the union of the records."""
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.metrics import create_metric
    g = torch.Generator().manual_seed(5)
    n = 4000
    uid = torch.randint(0, 60, (n,), generator=g)
    pred = torch.rand(n, generator=g)
    label = (torch.rand(n, generator=g) < pred).float()
    whole = create_metric("WuAucCalculator", name="w", uid_varname="uid")
    whole.add_data({"pred": pred, "label": label, "uid": uid})
    want = whole.get_metric_msg()
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({})
    box.init_metric("WuAucCalculator", "w", uid_varname="uid", metric_phase=1)
    mine = (uid % world) == rank
    box.get_metric("w").add_data({"pred": pred[mine], "label": label[mine], "uid": uid[mine]})
    got = box.get_wuauc_metric_msg("w")
    BoxWrapper.reset_instance()
    assert abs(got[4] - want[0]) < 1e-9 and abs(got[5] - want[1]) < 1e-9 and got[0] == want[2] and got[1] == want[3], (got, want)
    return got


def cpu_fleet_metrics_two_ranks(rank, world):
    """fleet.metrics over two workers = the single-process value on the union of the statistics."""
    from paddlebox_b200 import fleet
    g = torch.Generator().manual_seed(11)
    n, buckets = 3000, 256
    pred = torch.rand(n, generator=g)
    label = (torch.rand(n, generator=g) < pred).long()
    mine = (torch.arange(n) % world) == rank

    def stats(p, l):
        b = (p * buckets).long().clamp(max=buckets - 1)
        return (torch.bincount(b[l == 1], minlength=buckets).view(1, -1), torch.bincount(b[l == 0], minlength=buckets).view(1, -1))

    class _Solo:
        def all_reduce(self, v, mode="sum"):
            return v
    want = fleet.metrics.auc(*stats(pred, label), util=_Solo())
    got = fleet.metrics.auc(*stats(pred[mine], label[mine]))
    err = (pred - label).abs()
    mae = fleet.metrics.mae(err[mine].sum().view(1), torch.tensor([float(mine.sum())]))
    rmse = fleet.metrics.rmse((err[mine] ** 2).sum().view(1), [float(mine.sum())])
    assert abs(got - want) < 1e-12 and abs(mae - float(err.mean())) < 1e-6 and abs(rmse - float((err ** 2).mean().sqrt())) < 1e-6
    assert fleet.metrics.sum([rank + 1.0]).tolist() == [3.0] and fleet.metrics.max([rank]).tolist() == [1.0]
    assert fleet.metrics.min(np.array([rank, 5])).tolist() == [0.0, 5.0]
    assert abs(fleet.metrics.acc([3.0 * (rank + 1)], [10.0]) - 9.0 / 20.0) < 1e-12
    return got
