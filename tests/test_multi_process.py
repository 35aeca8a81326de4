"""Multi-process tests: 2 ranks over gloo on CPU (host-side logic), and — on a box with >= 2
GPUs — the real peer-memory path: CUDA-IPC wiring, cross-GPU barriers, sharded pull/push,
symmetric-memory collectives, data-parallel DeepFM.  (The reference tests distributed code the
same way: N local processes on 127.0.0.1, test_dist_base.py:865-1130.)"""
import socket

import pytest
import torch
import torch.multiprocessing as mp

from tests import mp_workers as W


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(fn, world, backend, timeout=300, args=()):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=W.run, args=(fn, r, world, port, backend, q, *args)) for r in range(world)]
    for p in procs:
        p.start()
    results = {}
    try:
        for _ in range(world):
            rank, status, out = q.get(timeout=timeout)
            results[rank] = (status, out)
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
    for r in range(world):
        assert r in results, f"rank {r} produced no result"
        assert results[r][0] == "ok", f"rank {r} failed:\n{results[r][1]}"
    return [results[r][1] for r in range(world)]


def test_two_ranks_gloo_cpu():
    outs = _spawn(W.cpu_feed_pass_and_dense, 2, "gloo")
    assert abs(outs[0] - outs[1]) < 1e-12 and 0.6 < outs[0] < 0.95     # same global AUC on both ranks


def test_global_shuffle_and_cluster_two_ranks_gloo():
    outs = _spawn(W.cpu_global_shuffle_and_cluster, 2, "gloo")
    assert outs[0]["before"] == 100 and outs[1]["before"] == 140
    assert outs[0]["total_before"] == outs[0]["total_after"] == 240          # nothing lost, nothing duplicated
    assert outs[0]["after"] + outs[1]["after"] == 240 and outs[0]["after"] != 100   # records really moved
    assert outs[0]["biggest"] == max(outs[0]["after"], outs[1]["after"])
    assert outs[0]["sums"] == [3.0, 2.0] and outs[1]["ranks"] == [0, 1]


def test_collective_ops_two_ranks_gloo():
    assert all(_spawn(W.cpu_collective_ops, 2, "gloo"))


def test_multi_node_dense_sync_four_ranks_gloo():
    outs = _spawn(W.cpu_multi_node_dense, 4, "gloo")
    assert outs[0] == outs[3] and set(outs[0]) == {"allreduce", "kstep", "sharding"}


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_sparse_over_all_to_all_gloo(world):
    outs = _spawn(W.cpu_sharded_sparse_router, world, "gloo")
    assert all(o[1] < o[0] for o in outs)


def test_model_saved_by_one_rank_loads_into_three_and_back(tmp_path):
    import numpy as np
    from paddlebox_b200.host import HostTable
    t = HostTable(4)
    keys = np.arange(1, 2001, dtype=np.uint64)
    rows = t.fetch_rows(keys)
    rows[:, 2] = (keys % 97).astype(np.float32)
    t.store_rows(keys, rows)
    ckpt = str(tmp_path / "one")
    assert t.save(ckpt) == 2000
    outs = _spawn(W.cpu_reshard_on_load, 3, "gloo", args=(ckpt,))
    assert outs == [2000, 2000, 2000]
    t2 = HostTable(4)                                              # 3 ranks -> 1 process again
    assert t2.load(ckpt + "_resaved_3") == 2000
    assert np.allclose(t2.fetch_rows(keys, create=False), t.fetch_rows(keys, create=False), rtol=1e-6)


def test_sparse_router_edge_cases_gloo():
    assert all(_spawn(W.cpu_router_edge_cases, 2, "gloo"))


def test_one_rank_per_node_topology_gloo():
    assert all(_spawn(W.cpu_one_rank_per_node, 3, "gloo"))


def test_train_from_dataset_trains_every_local_record_two_ranks(tmp_path):
    """with one process per GPU each rank loads files[rank::world]; the trainer must train
    all of them (not 1/world of them) and all ranks must step equally often."""
    from tests.util_data import write_files
    d = tmp_path / "data"
    d.mkdir()
    write_files(str(d), n_files=1, lines_per_file=200, n_sparse=3, n_dense=2, key_space=30)       # rank 0's file
    import os
    os.rename(d / os.listdir(d)[0], d / "a_part")
    write_files(str(d), n_files=1, lines_per_file=120, n_sparse=3, n_dense=2, key_space=30)       # rank 1's file
    other = [f for f in os.listdir(d) if f != "a_part"][0]
    os.rename(d / other, d / "b_part")
    outs = _spawn(W.cpu_train_from_dataset_sharded, 2, "gloo", args=(str(d),))
    assert outs[0]["loaded"] == 200 and outs[1]["loaded"] == 120
    assert outs[0]["instances"] == 200 and outs[1]["instances"] == 120           # nothing dropped
    assert outs[0]["steps"] == outs[1]["steps"] == 13                               # ceil(200 / 16), equalised


def test_overlapped_bucketed_dense_allreduce_two_ranks_gloo():
    outs = _spawn(W.cpu_dense_overlap, 2, "gloo")
    assert outs[0]["buckets"] >= 3 and outs[0] == outs[1]


def test_psgpu_pre_build_thread_has_its_own_group_two_ranks_gloo():
    outs = _spawn(W.cpu_psgpu_pipeline_two_ranks, 2, "gloo")
    assert all(len(o) == 3 and min(o) > 0 for o in outs)


def test_queue_dataset_streams_on_two_ranks_with_unequal_shares(tmp_path):
    from tests.util_data import write_files
    write_files(str(tmp_path), n_files=5, lines_per_file=40, n_sparse=3, n_dense=2, key_space=60)
    outs = _spawn(W.cpu_queue_dataset_streaming_two_ranks, 2, "gloo", args=(str(tmp_path),))
    # rank 1 has two files of 40 lines = 8 batches of 10: both ranks stop there
    assert outs[0]["steps"] == outs[1]["steps"] == 8 and outs[1]["instances"] == 80


def test_user_sharded_file_lists_train_everything_and_metrics_are_global(tmp_path):
    from tests.util_data import write_files
    write_files(str(tmp_path), n_files=4, lines_per_file=40, n_sparse=3, n_dense=2, key_space=60)
    outs = _spawn(W.cpu_user_sharded_files_and_global_metrics, 2, "gloo", args=(str(tmp_path),))
    for o in outs:
        assert o["local"] == 80 and o["trained"] == 80, o          # every local record, not 1 / world of them
        assert o["metric_size"] == 160, o                          # reduced over both ranks
    assert outs[0]["steps"] == outs[1]["steps"] == 8


def test_next_pass_is_fed_while_the_current_one_trains_two_ranks(tmp_path):
    outs = _spawn(W.cpu_preload_overlap_two_ranks, 2, "gloo", args=(str(tmp_path),))
    assert outs[0] == outs[1]
    # 4 files per pass of 48 / 64 / 48 lines: the metric covers both ranks' records
    assert outs[0]["sizes"] == [4 * 48, 4 * 64, 4 * 48] and outs[0]["keys_total"] > 150


def test_fleet_metrics_reduce_statistics_over_workers():
    outs = _spawn(W.cpu_fleet_metrics_two_ranks, 2, "gloo")
    assert outs[0] == outs[1] and 0.5 < outs[0] < 1.0


def test_per_user_auc_is_reduced_over_ranks():
    outs = _spawn(W.cpu_wuauc_two_ranks, 2, "gloo")
    assert outs[0] == outs[1] and outs[0][0] == 60 and outs[0][1] == 4000


def test_legacy_record_global_shuffle_two_ranks_gloo():
    outs = _spawn(W.cpu_legacy_global_shuffle, 2, "gloo")
    assert outs[0]["total_before"] == outs[0]["total_after"] == 110 and all(o["owner_ok"] for o in outs)
    assert outs[0]["after"] + outs[1]["after"] == 110 and outs[0]["after"] not in (0, 110)
    assert outs[0]["sum_signs"] + outs[1]["sum_signs"] == sum(range(50)) + sum(1000 + i for i in range(60))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.gpu
@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sparse_and_collectives_real_gpus(world):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    assert all(_spawn(W.gpu_sparse_and_dense, world, "nccl"))


@pytest.mark.gpu
@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_deepfm_data_parallel_real_gpus():
    outs = _spawn(W.gpu_deepfm_data_parallel, 2, "nccl")
    assert outs[0][1] < outs[0][0]
