"""Pass build / write-back pipeline (pass_build.py): chunking, slot recycling, completeness, error
propagation — on the CPU with the real native host table and a recording device stage."""
import threading
import time

import numpy as np
import pytest
import torch

from paddlebox_b200.host import HostTable
from paddlebox_b200.pass_build import PassBuilder, StagingPool


def test_build_streams_every_row_once_through_a_small_pool():
    t = HostTable(4)
    keys = np.arange(1, 10_001, dtype=np.uint64)
    rows = t.fetch_rows(keys)
    rows[:, 2] = (keys % 89).astype(np.float32)
    t.store_rows(keys, rows)
    got_k, got_r, seen_slots = [], [], set()

    def insert(dk, dr):
        seen_slots.add(dk.data_ptr())
        got_k.append(dk.clone())
        got_r.append(dr.clone())
        time.sleep(0.001)                         # the device stage is slower than the fetch: slots must recycle
    pb = PassBuilder(t, insert, t.layout.stride, None, chunk=777, depth=3, fetch_threads=3)
    st = pb.build(keys)
    k = torch.cat(got_k).numpy().view(np.uint64)
    r = torch.cat(got_r).numpy()
    assert st["rows"] == 10_000 and st["chunks"] == 13 and len(seen_slots) <= 3
    order = np.argsort(k)
    assert np.array_equal(k[order], keys) and np.array_equal(r[order], t.fetch_rows(keys, create=False))
    assert st["rows_per_s"] > 0 and st["staging_host_bytes"] == 3 * 777 * (8 + 4 * t.layout.stride)
    # keys the host tier has never seen are created on the way (BeginPass semantics)
    new = np.arange(50_000, 50_100, dtype=np.uint64)
    got_k.clear(); got_r.clear()
    pb.build(new)
    assert t.size() == 10_100 and sum(x.numel() for x in got_k) == 100


def test_write_back_in_bounded_chunks_and_error_propagation():
    t = HostTable(4)
    S = t.layout.stride
    keys = torch.arange(1, 5001, dtype=torch.int64)
    rows = torch.rand(5000, S)
    rows[:, S - 1] = 0
    pb = PassBuilder(t, lambda *_: None, S, None, chunk=512, depth=2)
    st = pb.write_back(keys, rows)
    assert st["rows"] == 5000 and t.size() == 5000
    back = t.fetch_rows(keys.numpy().view(np.uint64), create=False)
    assert np.allclose(back[:, :S - 1], rows.numpy()[:, :S - 1])

    class Broken:
        layout = t.layout

        def fetch_rows(self, keys, create=True):
            raise RuntimeError("host tier unavailable")
    pb2 = PassBuilder(Broken(), lambda *_: None, S, None, chunk=100, depth=2)
    with pytest.raises(RuntimeError, match="host tier unavailable"):
        pb2.build(np.arange(1, 1000, dtype=np.uint64))
    assert threading.active_count() < 20          # the fetch workers were released


def test_staging_pool_geometry():
    p = StagingPool(1000, 32, None, depth=4)
    assert len(p.slots) == 4 and p.bytes_per_row() == 8 + 128 and p.host_bytes() == 4 * 1000 * 136


def test_concurrent_bulk_calls_on_one_host_table():
    """The pass builder's fetch threads, and a write-back running next to the next pass's BuildPull, call the
    bulk operations of ONE table at the same time (each call is itself parallel over the table's shards):
    every key must end up in the table once, with the row its writer stored."""
    import threading
    import numpy as np
    from paddlebox_b200.host import HostTable
    t = HostTable(embedx_dim=8)
    stride = t.layout.stride
    rng = np.random.default_rng(0)
    keys = np.unique(rng.integers(1, 1 << 48, 400_000, dtype=np.uint64))
    parts = np.array_split(keys, 8)
    errs = []

    def creator(k):
        try:
            rows = t.fetch_rows(k, create=True)
            assert rows.shape == (k.size, stride)
        except BaseException as e:  # noqa: BLE001
            errs.append(e)

    def writer(k, tag):
        try:
            rows = np.full((k.size, stride), tag, np.float32)
            t.store_rows(k, rows)
            t.pull(k)
        except BaseException as e:  # noqa: BLE001
            errs.append(e)

    ths = [threading.Thread(target=creator, args=(p,)) for p in parts]
    ths += [threading.Thread(target=writer, args=(parts[i], float(i + 1))) for i in range(0, 8, 2)]
    ths += [threading.Thread(target=creator, args=(keys[::3],))]              # overlaps everybody's shards
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    assert not errs, errs
    assert t.size() == keys.size
    assert np.array_equal(np.sort(t.keys()), keys)
    for i in range(0, 8, 2):
        got = t.fetch_rows(parts[i], create=False)
        # the writer's rows, unless that key's creator ran after it (then the row is still the writer's: creation
        # never overwrites an existing row)
        assert np.all(got[:, 0] == float(i + 1))


def test_pass_builder_with_several_fetch_threads_at_scale():
    """More than one chunk AND more than one fetch thread (the default configuration on a real pass)."""
    import numpy as np
    from paddlebox_b200.host import HostTable
    from paddlebox_b200.pass_build import PassBuilder
    t = HostTable(embedx_dim=8)
    keys = np.unique(np.random.default_rng(1).integers(1, 1 << 48, 300_000, dtype=np.uint64))
    seen = []
    pb = PassBuilder(t, lambda k, r: seen.append(k.numpy().copy().view(np.uint64)), t.layout.stride, None,
                     chunk=1 << 14, depth=3, fetch_threads=4)
    r = pb.build(keys)
    assert r["chunks"] == (keys.size + (1 << 14) - 1) // (1 << 14) and t.size() == keys.size
    assert np.array_equal(np.sort(np.concatenate(seen)), keys)
