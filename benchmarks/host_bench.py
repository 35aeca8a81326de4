#!/usr/bin/env python
"""Host-side throughput of the data path (no GPU needed): text parse + load, batch packing into
pinned-style staging buffers, host-tier pull/push, AUC accumulation.

    python benchmarks/host_bench.py [--instances 200000] [--json out.json]

The GPU trains a 4096-instance DeepFM batch in ~0.3 ms, i.e. ~3300 batches/s per GPU: the
packer numbers say how many loader threads a real (non-synthetic) feed needs per GPU."""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--instances", type=int, default=200_000)
    ap.add_argument("--slots", type=int, default=26)
    ap.add_argument("--dense", type=int, default=13)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.host import HostAuc, HostTable
    from tests.util_data import slot_vars, write_files
    out = {"cpu_count": os.cpu_count(), "instances": a.instances, "slots": a.slots, "batch": a.batch}
    with tempfile.TemporaryDirectory() as tmp:
        nfiles = max(1, a.threads)
        t0 = time.time()
        files = write_files(tmp, n_files=nfiles, lines_per_file=a.instances // nfiles, n_sparse=a.slots,
                            n_dense=a.dense, key_space=1_000_000, max_len=5)
        out["write_files_s"] = round(time.time() - t0, 2)
        nbytes = sum(os.path.getsize(f) for f in files)
        ds = DatasetFactory().create_dataset("InMemoryDataset")
        ds.set_use_var(slot_vars(a.slots, a.dense))
        ds.set_batch_size(a.batch)
        ds.set_thread(a.threads)
        ds.set_filelist(files)
        t0 = time.time()
        ds.load_into_memory()
        dt = time.time() - t0
        n = ds.get_memory_data_size()
        out["load"] = {"seconds": round(dt, 3), "instances_per_s": round(n / dt), "MB_per_s": round(nbytes / dt / 1e6, 1),
                       "threads": a.threads}
        t0 = time.time()
        ds.prepare_train(1)
        out["prepare_train_s"] = round(time.time() - t0, 3)
        # packing: host_batches() = pack into staging buffers, no device work
        for rep in range(2):
            t0 = time.time()
            nb = nk = 0
            for hb in ds.host_batches():
                nb += 1
                nk += hb.n_keys
                hb.release()
            dt = time.time() - t0
        out["pack"] = {"batches_per_s": round(nb / dt, 1), "keys_per_s": round(nk / dt), "batches": nb,
                       "keys_per_batch": round(nk / max(nb, 1))}
        # pass file: the prepared pass written as-is and mapped back (no decode, no key collection)
        pf = os.path.join(tmp, "pass.arena")
        t0 = time.time()
        ds.save_pass_file(pf)
        t_save = time.time() - t0
        ds2 = DatasetFactory().create_dataset("InMemoryDataset")
        ds2.set_use_var(slot_vars(a.slots, a.dense))
        ds2.set_batch_size(a.batch)
        t0 = time.time()
        ds2.load_pass_file(pf)
        ds2.prepare_train(1)
        t_load = time.time() - t0
        t0 = time.time()
        nb2 = 0
        for hb in ds2.host_batches():
            nb2 += 1
            hb.release()
        t_pack = time.time() - t0
        out["pass_file"] = {"MB": round(os.path.getsize(pf) / 1e6, 1), "save_instances_per_s": round(n / t_save),
                            "load_and_prepare_instances_per_s": round(n / t_load),
                            "pack_batches_per_s_from_mapping": round(nb2 / max(t_pack, 1e-9), 1)}
        ds2.release_memory()
    # host tier
    tab = HostTable(embedx_dim=16)
    keys = np.unique(np.random.default_rng(0).integers(1, 1 << 40, 1_000_000, dtype=np.uint64))
    t0 = time.time()
    tab.fetch_rows(keys, create=True)
    out["host_table_create_rows_per_s"] = round(keys.size / (time.time() - t0))
    t0 = time.time()
    tab.pull(keys)
    out["host_table_pull_rows_per_s"] = round(keys.size / (time.time() - t0))
    push = torch.zeros(keys.size, tab.layout.push_stride)
    push[:, 1] = 1.0
    push[:, 3:] = 0.01
    t0 = time.time()
    tab.push(keys, push)
    out["host_table_push_rows_per_s"] = round(keys.size / (time.time() - t0))
    # end of a feed pass: sort + unique of the pass keys, and their partition by owning rank
    from paddlebox_b200.host import partition_by_owner, sort_unique_u64
    raw = np.random.default_rng(1).integers(1, 1 << 62, 20_000_000, dtype=np.uint64)
    t0 = time.time()
    uk = sort_unique_u64(raw)
    out["sort_unique_keys_per_s"] = round(raw.size / (time.time() - t0))
    t0 = time.time()
    partition_by_owner(uk, 8)
    out["partition_by_owner_keys_per_s"] = round(uk.size / (time.time() - t0))
    # SSD tier: spill half of the table, then promote it back
    import tempfile as _tf
    with _tf.TemporaryDirectory() as ssd:
        tab.ssd_open(ssd)
        tab.save(os.path.join(ssd, "delta"), mode=1, clear_dirty=True)      # clean rows may spill
        t0 = time.time()
        n_sp = tab.spill(max_mem_rows=keys.size // 2)
        out["ssd_spill_rows_per_s"] = round(n_sp / max(time.time() - t0, 1e-9))
        t0 = time.time()
        n_ld = tab.load_ssd2mem()
        out["ssd_load_rows_per_s"] = round(n_ld / max(time.time() - t0, 1e-9))
        tab.ssd_close(True)
    auc = HostAuc(1_000_000)
    p, y = np.random.rand(4_000_000).astype(np.float32), (np.random.rand(4_000_000) < 0.2).astype(np.float32)
    t0 = time.time()
    auc.add(p, y)
    out["auc_add_instances_per_s"] = round(p.size / (time.time() - t0))
    print(json.dumps(out, indent=1))
    if a.json:
        json.dump(out, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
