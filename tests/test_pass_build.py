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
