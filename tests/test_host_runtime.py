"""Host C++ runtime: parser, batch dealing, dataset, host table, AUC, checkpoints (CPU)."""
import math
import os

import numpy as np
import pytest
import torch

from paddlebox_b200 import _native as N
from paddlebox_b200.host import HostAuc, HostTable, NativeDataset, compute_batch_ranges, AsyncDenseTable
from paddlebox_b200.ops import reference as R
from tests.util_data import slot_vars, write_files


# ---------------------------------------------------------------- parser
def _ds(vars_):
    d = NativeDataset()
    d.set_slots([v.desc() for v in vars_])
    return d


def test_parse_line_basic_and_drops_zero_feasigns():
    from paddlebox_b200.dataset import dense_slot, sparse_slot
    d = _ds([dense_slot("label", 1), sparse_slot("a"), sparse_slot("b")])
    u, f, _ = d.parse_line("1 1.0 3 11 0 13 1 7")
    assert u.tolist() == [11, 13, 7]          # zero feasign dropped (data_feed.cc:4099)
    assert f.tolist() == [1.0]
    assert d.parse_line("1 1.0 1 0 1 0") is None          # no feasign at all -> rejected
    assert d.parse_line("1 1.0 0 1 7") is None             # n == 0 is illegal
    assert d.parse_line("1 1.0 2 5") is None                # truncated


def test_parse_line_ins_id_logkey_and_unused_slot():
    from paddlebox_b200.data_generator import make_logkey
    from paddlebox_b200.dataset import dense_slot, sparse_slot, SlotVar
    d = NativeDataset()
    slots = [dense_slot("label", 1), SlotVar("skipme", "int64", 1, (-1, 1), is_used=False), sparse_slot("a")]
    d.set_slots([s.desc() for s in slots])
    d.set("parse_ins_id", 1)
    d.set("parse_logkey", 1)
    lk = make_logkey(0xabcdef12345, 0x0de, 2)
    u, f, sid = d.parse_line(f"1 ins_42 1 {lk} 1 0.0 2 99 98 1 5")
    assert u.tolist() == [5] and sid == 0xabcdef12345


# ---------------------------------------------------------------- batch dealing
@pytest.mark.parametrize("n,bs,w", [(1000, 64, 4), (8 * 64 * 3 + 5, 64, 8), (100, 64, 8), (7, 4, 8), (4096 * 9, 4096, 8),
                                    (129, 16, 2)])
def test_batch_ranges_equal_counts(n, bs, w):
    r = compute_batch_ranges(n, bs, w)
    assert sum(l for _, l in r) == n
    assert len(r) % w == 0                         # every worker gets the same number of batches
    pos = 0
    for s, l in r:
        assert s == pos and l >= 0
        pos += l
    assert max(l for _, l in r) <= max(bs, n // w + n % w) * 2


# ---------------------------------------------------------------- dataset
def test_dataset_load_shuffle_batches(tmp_path):
    vars_ = slot_vars(6, 3)
    files = write_files(str(tmp_path), n_files=3, lines_per_file=101, n_sparse=6, n_dense=3, key_space=50)
    d = _ds(vars_)
    d.set_filelist(files)
    d.set("thread_num", 2)
    d.set("batch_size", 32)
    d.load_into_memory()
    assert d.num_records() == 303
    uk = d.unique_keys()
    assert uk[0] == 0 and len(np.unique(uk)) == len(uk)   # key 0 always registered
    d.prepare_train(2, True, 7, 0)
    assert d.num_batches(0) == d.num_batches(1) > 0
    seen_keys = set()
    total = 0
    for w in range(2):
        for b in range(d.num_batches(w)):
            B, n = d.batch_shape(w, b)
            keys = np.zeros(n, dtype=np.uint64)
            off = np.zeros(6 * B + 1, dtype=np.int32)
            dense = np.zeros((B, 4), dtype=np.float32)
            d.pack_batch(w, b, keys.ctypes.data, off.ctypes.data, dense.ctypes.data, True)
            assert off[-1] == n and np.all(np.diff(off) >= 1)
            assert set(np.unique(dense[:, 0])) <= {0.0, 1.0}
            seen_keys |= set(keys.tolist())
            total += B
            # record-major packing holds the same multiset of keys
            keys2 = np.zeros(n, dtype=np.uint64)
            off2 = np.zeros(B * 7, dtype=np.int32)
            d.pack_batch(w, b, keys2.ctypes.data, off2.ctypes.data, dense.ctypes.data, False)
            assert sorted(keys2.tolist()) == sorted(keys.tolist())
    assert total == 303
    assert seen_keys | {0} == set(uk.tolist())


def test_dataset_pipe_command_sample_rate_archive(tmp_path):
    vars_ = slot_vars(3, 0)
    files = write_files(str(tmp_path), n_files=1, lines_per_file=400, n_sparse=3, n_dense=0, key_space=30)
    d = _ds(vars_)
    d.set_filelist(files)
    d.set("pipe_command", "head -n 100")
    d.load_into_memory()
    assert d.num_records() == 100
    arch = str(tmp_path / "arch.bin")
    d.save_archive(arch)
    keys_before = d.unique_keys()
    d2 = _ds(vars_)
    d2.load_archive(arch)
    assert d2.num_records() == 100 and np.array_equal(d2.unique_keys(), keys_before)
    d3 = _ds(vars_)
    d3.set_filelist(files)
    d3.set("sample_rate", 0.5)
    d3.load_into_memory()
    assert 120 < d3.num_records() < 280


def test_dataset_preload_and_slots_shuffle(tmp_path):
    vars_ = slot_vars(4, 0)
    files = write_files(str(tmp_path), n_files=2, lines_per_file=50, n_sparse=4, n_dense=0, key_space=1000)
    d = _ds(vars_)
    d.set_filelist(files)
    d.preload_into_memory()
    d.wait_preload_done()
    assert d.num_records() == 100
    d.prepare_train(1, False, 0, 0)

    def slot_keys(slot):
        out = []
        for b in range(d.num_batches(0)):
            B, n = d.batch_shape(0, b)
            keys = np.zeros(n, dtype=np.uint64)
            off = np.zeros(4 * B + 1, dtype=np.int32)
            d.pack_batch(0, b, keys.ctypes.data, off.ctypes.data, 0, True)
            out.append(keys[off[slot * B]:off[(slot + 1) * B]].tolist())
        return out
    before0, before1 = slot_keys(1), slot_keys(0)
    d.slots_shuffle([2], 3)     # index 2 in the slot list = sparse slot1 (label is slot 0)
    after0, after1 = slot_keys(1), slot_keys(0)
    assert after1 == before1                              # untouched slot
    assert sorted(sum(after0, [])) == sorted(sum(before0, [])) and after0 != before0


def test_parser_plugin(tmp_path):
    """A user parser .so (role of ISlotParser plugins, data_feed.cc:3604-3670)."""
    import subprocess
    src = tmp_path / "plug.cc"
    src.write_text(r'''
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
extern "C" int pbx_parse_line(const char* line, size_t len, void* ctx,
    void (*emit_u64)(void*, int, const uint64_t*, size_t), void (*emit_f32)(void*, int, const float*, size_t),
    void (*emit_meta)(void*, const char*, uint64_t, uint32_t, uint32_t)) {
  // format: "<label>;<k1>,<k2>,..."
  char* end = nullptr;
  float label = strtof(line, &end);
  if (*end != ';') return 1;
  std::vector<uint64_t> ks;
  const char* p = end + 1;
  while (p < line + len) { ks.push_back(strtoull(p, &end, 10)); if (end == p) break; p = end + 1; }
  emit_f32(ctx, 0, &label, 1);
  emit_u64(ctx, 1, ks.data(), ks.size());
  emit_meta(ctx, "x", 7, 0, 0);
  return 0;
}''')
    so = tmp_path / "plug.so"
    subprocess.check_call(["g++", "-shared", "-fPIC", "-O1", "-o", str(so), str(src)])
    data = tmp_path / "d.txt"
    data.write_text("1;5,6,7\n0;8\n")
    from paddlebox_b200.dataset import dense_slot, sparse_slot
    d = NativeDataset()
    d.set("parser_plugin", str(so))
    d.set_slots([dense_slot("label").desc(), sparse_slot("a").desc()])
    d.set_filelist([str(data)])
    d.load_into_memory()
    assert d.num_records() == 2 and set(d.unique_keys().tolist()) == {0, 5, 6, 7, 8}


def test_parser_plugin_whole_file_mode(tmp_path):
    """pbx_parse_file: the plugin reads a binary container itself and reports record after record
    (FLAGS_enable_ins_parser_file)."""
    import struct
    import subprocess
    src = tmp_path / "plugf.cc"
    src.write_text(r'''
#include <cstdint>
#include <cstdio>
#include <vector>
// file = records of: float label, uint32 n, n x uint64 keys
extern "C" int pbx_parse_file(const char* path, void* ctx, void (*begin_record)(void*),
    void (*emit_u64)(void*, int, const uint64_t*, size_t), void (*emit_f32)(void*, int, const float*, size_t),
    void (*emit_meta)(void*, const char*, uint64_t, uint32_t, uint32_t), int (*end_record)(void*)) {
  FILE* fp = fopen(path, "rb");
  if (!fp) return 1;
  float label; uint32_t n; uint64_t sid = 0;
  while (fread(&label, 4, 1, fp) == 1 && fread(&n, 4, 1, fp) == 1) {
    std::vector<uint64_t> ks(n);
    if (n && fread(ks.data(), 8, n, fp) != n) { fclose(fp); return 2; }
    begin_record(ctx);
    emit_f32(ctx, 0, &label, 1);
    emit_u64(ctx, 1, ks.data(), ks.size());
    emit_meta(ctx, "f", ++sid, 0, 0);
    end_record(ctx);
  }
  fclose(fp);
  return 0;
}''')
    so = tmp_path / "plugf.so"
    subprocess.check_call(["g++", "-shared", "-fPIC", "-O1", "-o", str(so), str(src)])
    recs = [(1.0, [5, 6, 7]), (0.0, [8]), (1.0, []), (0.0, [0, 9])]     # third has no feasign: rejected
    for name in ("a.bin", "b.bin"):
        with open(tmp_path / name, "wb") as f:
            for label, ks in recs:
                f.write(struct.pack("<fI", label, len(ks)) + struct.pack(f"<{len(ks)}Q", *ks))
    from paddlebox_b200.dataset import dense_slot, sparse_slot
    d = NativeDataset()
    d.set("parser_plugin", str(so))
    d.set("parser_file_mode", 1)
    d.set_slots([dense_slot("label").desc(), sparse_slot("a").desc()])
    d.set("thread_num", 2)
    d.set_filelist([str(tmp_path / "a.bin"), str(tmp_path / "b.bin")])
    d.load_into_memory()
    assert d.num_records() == 6 and set(d.unique_keys().tolist()) == {0, 5, 6, 7, 8, 9}
    assert "records=6" in d.stats() and "bad_lines=2" in d.stats()


# ---------------------------------------------------------------- host table
@pytest.mark.parametrize("opt", ["adagrad", "adam", "shared_adam"])
def test_host_table_matches_python_oracle(opt):
    cfg = N.SparseOptConfig(mf_create_thresholds=2.0, mf_initial_range=0.0)
    t = HostTable(4, opt, cfg)
    keys = np.arange(1, 40, dtype=np.uint64) * 13
    t.fetch_rows(keys, create=True)
    rows = {int(k): R.new_row(t.layout, cfg) for k in keys}
    g = torch.Generator().manual_seed(0)
    for _ in range(5):
        push = torch.zeros(len(keys), t.layout.push_stride)
        push[:, 0] = 2
        push[:, 1] = torch.randint(1, 4, (len(keys),), generator=g).float()
        push[:, 2] = torch.randint(0, 2, (len(keys),), generator=g).float()
        push[:, 3:8] = torch.randn(len(keys), 5, generator=g)
        t.push(keys, push)
        for i, k in enumerate(keys):
            rows[int(k)] = R.sparse_row_update(rows[int(k)], push[i].tolist(), t.layout, cfg)
    got = t.fetch_rows(keys, create=False)
    exp = np.array([rows[int(k)] for k in keys], dtype=np.float32)
    assert np.allclose(got, exp, rtol=2e-4, atol=2e-5)
    pulled = t.pull(np.concatenate([keys[:3], np.array([0, 999999], dtype=np.uint64)]))
    assert torch.allclose(pulled[:3], torch.from_numpy(exp[:3, :7]), atol=1e-6) and torch.all(pulled[3:] == 0)


def test_host_table_checkpoint_roundtrip_and_shrink(tmp_path):
    cfg = N.SparseOptConfig(mf_create_thresholds=1.0)
    t = HostTable(4, "adagrad", cfg)
    keys = np.arange(1, 101, dtype=np.uint64)
    t.fetch_rows(keys)
    push = np.zeros((100, t.layout.push_stride), dtype=np.float32)
    push[:, 1] = 30
    push[:, 2] = 5
    push[:, 3:8] = np.random.RandomState(0).randn(100, 5)
    t.push(keys[:60], push[:60])
    t.push(keys[:60], push[:60])
    n = t.save(str(tmp_path / "batch"), mode=0)
    assert n == 100
    files = sorted(os.listdir(tmp_path / "batch"))
    assert files[0].startswith("part-000-") and len(files) == 64
    t2 = HostTable(4, "adagrad", cfg)
    assert t2.load(str(tmp_path / "batch")) == 100
    assert np.allclose(t2.fetch_rows(keys, create=False), t.fetch_rows(keys, create=False), rtol=1e-6, atol=1e-7)
    # delta: only touched keys, serving format (weights only)
    nd = t.save(str(tmp_path / "delta"), mode=1)
    assert nd == 60
    assert t.save(str(tmp_path / "delta2"), mode=1) == 0      # dirty flags were cleared
    t3 = HostTable(4, "adagrad", cfg)
    assert t3.load(str(tmp_path / "delta"), serving=True) == 60
    a, b = t3.pull(keys[:60]), t.pull(keys[:60])
    assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)
    kept = t.shrink(decay=0.5, delete_threshold=1.0, delete_after_unseen_days=30)
    assert kept == 60 and t.size() == 60


def test_host_table_ssd_tier(tmp_path):
    """Cold rows spill to the per-shard SSD files, come back on lookup / push / LoadSSD2Mem, and are
    part of every full save; dirty rows stay in memory."""
    cfg = N.SparseOptConfig(mf_create_thresholds=0.0, mf_initial_range=0.01)
    t = HostTable(4, "adagrad", cfg)
    t.ssd_open(str(tmp_path / "ssd"))
    keys = np.arange(1, 5001, dtype=np.uint64)
    rows = t.fetch_rows(keys)
    rows[:, 2] = (keys % 97).astype(np.float32)
    age = t.layout.off_tail + 2
    rows[:, age] = (keys % 10).astype(np.float32)           # unseen days 0..9
    t.store_rows(keys, rows, mark_dirty=False)
    t.store_rows(keys[:50], rows[:50], mark_dirty=True)      # touched in the current delta window
    ref = t.fetch_rows(keys, create=False).copy()
    # by age: every clean row unseen for >= 5 days goes
    n_old = int(((keys % 10) >= 5)[50:].sum())
    assert t.spill(min_unseen_days=5.0) == n_old
    assert t.ssd_rows() == n_old and t.mem_rows() == 5000 - n_old and t.size() == 5000
    # by budget: the memory tier shrinks to about the limit, oldest first; dirty rows stay
    t.spill(max_mem_rows=1280)
    assert t.mem_rows() <= 1280 + 64 + 50 and t.size() == 5000
    assert len(np.unique(t.keys())) == 5000
    # a full save streams the cold rows too; a delta holds the dirty rows only
    assert t.save(str(tmp_path / "batch"), mode=0, clear_dirty=False) == 5000
    assert t.save(str(tmp_path / "delta"), mode=1, clear_dirty=False) == 50
    t2 = HostTable(4, "adagrad", cfg)
    assert t2.load(str(tmp_path / "batch")) == 5000
    assert np.allclose(t2.fetch_rows(keys, create=False), ref, rtol=1e-6, atol=1e-7)
    # pull promotes exactly what it touches, values intact
    cold_before = t.ssd_rows()
    probe = keys[4000:4100]
    got = t.pull(probe)
    assert torch.equal(got[:, 2], torch.from_numpy((probe % 97).astype(np.float32)))
    assert t.size() == 5000 and t.ssd_rows() <= cold_before
    # a push on a cold key updates the promoted row
    push = np.zeros((1, t.layout.push_stride), dtype=np.float32)
    push[0, 1] = 3
    k = keys[4999:5000]
    show0 = ref[4999, 0]
    t.push(k, push)
    assert t.fetch_rows(k, create=False)[0, 0] == show0 + 3
    # spill / promote cycles leave garbage behind; the files are compacted and nothing is lost
    for _ in range(4):
        t.save(str(tmp_path / "d"), mode=1, clear_dirty=True)
        t.spill(max_mem_rows=640)
        t.pull(keys[::3])
    assert t.size() == 5000
    # LoadSSD2Mem brings everything back; rows equal the reference except the pushed one
    n_cold = t.ssd_rows()
    assert t.load_ssd2mem() == n_cold and t.ssd_rows() == 0 and t.mem_rows() == 5000
    back = t.fetch_rows(keys, create=False)
    assert np.allclose(back[:4999], ref[:4999], rtol=1e-6, atol=1e-7)
    # shrink sees cold rows as well
    t.spill(max_mem_rows=640)
    kept = t.shrink(decay=1.0, delete_threshold=-1e9, delete_after_unseen_days=7.5)
    assert kept == 3500 and t.size() == 3500 and t.ssd_rows() == 0      # ages 0..6 survive one more day
    t.ssd_close(unlink=True)
    assert not any(f.endswith(".bin") for f in os.listdir(tmp_path / "ssd"))


def test_box_wrapper_spills_at_end_pass(tmp_path):
    from paddlebox_b200.box_wrapper import BoxWrapper
    BoxWrapper.reset_instance()
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    try:
        box.initialize_gpu_and_load_model({"ssd_dir": str(tmp_path / "ssd"), "host_mem_rows_limit": 640})
        for day, lo in (("20240101", 1), ("20240102", 3001)):
            box.begin_feed_pass(day)
            box.feed_pass(np.arange(lo, lo + 3000, dtype=np.uint64))
            box.end_feed_pass()
            box.begin_pass()
            box.end_pass()
            box.save_delta(str(tmp_path / f"delta{day}"))
        assert box.check_need_limit_mem() is False or box.host_table.mem_rows() <= 3000 + 640
        assert box.host_table.ssd_rows() > 0 and box.host_table.size() == 6001      # + key 0
        vals, restore = box.pull_sparse(torch.arange(1, 50, dtype=torch.int64))   # day-1 keys come back
        assert vals.shape[0] == 49
        assert box.load_ssd2mem("20240103") > 0 and box.host_table.ssd_rows() == 0
    finally:
        BoxWrapper.reset_instance()


# ---------------------------------------------------------------- metrics
def test_auc_matches_exact_and_components():
    rs = np.random.RandomState(0)
    pred = rs.rand(5000).astype(np.float32)
    label = (rs.rand(5000) < pred).astype(np.float32)
    a = HostAuc(100000)
    a.add(pred[:2500], label[:2500])
    a.add(pred[2500:], label[2500:])
    auc, bucket_err, mae, rmse, actr, pctr, copc, size = a.compute()
    # exact AUC by rank statistic
    order = np.argsort(pred)
    ranks = np.empty(5000)
    ranks[order] = np.arange(1, 5001)
    npos = label.sum()
    exact = (ranks[label == 1].sum() - npos * (npos + 1) / 2) / (npos * (5000 - npos))
    assert abs(auc - exact) < 2e-4
    assert size == 5000 and abs(actr - label.mean()) < 1e-9 and abs(pctr - pred.mean()) < 1e-6
    assert abs(mae - np.abs(pred - label).mean()) < 1e-6
    assert abs(rmse - math.sqrt(((pred - label) ** 2).mean())) < 1e-6
    assert abs(copc - actr / pctr) < 1e-9 and bucket_err >= 0
    b = HostAuc(1000)
    b.add(np.full(10, 0.3, np.float32), np.zeros(10, np.float32))
    assert b.compute()[0] == -0.5                 # all one class (metrics.cc:323-325)
    m = HostAuc(1000)
    m.add(pred, label, mask=(np.arange(5000) % 2).astype(np.float32))
    assert m.compute()[7] == 2500


def test_wuauc():
    a = HostAuc(1000)
    pred = np.array([0.9, 0.1, 0.8, 0.2, 0.5, 0.6], np.float32)
    label = np.array([1, 0, 0, 1, 1, 1], np.float32)
    uid = np.array([1, 1, 2, 2, 3, 3], np.uint64)
    a.add_uid(pred, label, uid)
    uauc, wuauc, users, size = a.compute_wuauc()
    assert users == 2 and size == 4            # user 3 has one class only
    assert abs(uauc - 0.5) < 1e-6 and abs(wuauc - 0.5) < 1e-6


def test_async_dense_table():
    t = AsyncDenseTable(1000, lr=0.1, threads=2, merge_limit=4)
    t.set_params(torch.ones(1000))
    for _ in range(8):
        t.push(torch.full((1000,), 0.5))
    t.wait()
    p = t.pull()
    assert t.updates() >= 2 and torch.all(p < 1.0) and torch.allclose(p, p[0].expand(1000))


def test_pv_batches_carry_rank_offset(tmp_path):
    """logkey-parsed, PV-merged batches expose search_id / rank / cmatch and the rank-offset matrix."""
    import torch
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.ops.ctr_ops import build_rank_offset
    from tests.util_data import slot_vars, write_files
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var(slot_vars(3, 2))
    ds.set_batch_size(12)
    ds.set_thread(1)
    ds.set_parse_logkey(True)
    ds.set_enable_pv_merge(True)
    ds.set_rank_offset("rank_offset")
    ds.set_filelist(write_files(str(tmp_path), n_files=1, lines_per_file=36, n_sparse=3, n_dense=2, key_space=30,
                                logkey=True))
    ds.load_into_memory()
    ds.prepare_train(1)
    seen = 0
    for b in ds.batches(device=torch.device("cpu")):
        m = b.meta
        assert m is not None and m["rank_offset"].shape == (b.batch_size, 7)
        sid, rk, ro = m["search_id"].tolist(), m["rank"].tolist(), m["rank_offset"]
        # brute-force definition
        for i in range(b.batch_size):
            if not 1 <= rk[i] <= 3:
                assert int(ro[i, 0]) == -1
                continue
            assert int(ro[i, 0]) == rk[i]
            for k in range(3):
                js = [j for j in range(b.batch_size) if sid[j] == sid[i] and rk[j] == k + 1]
                if js:
                    assert int(ro[i, 2 * k + 1]) == k + 1 and int(ro[i, 2 * k + 2]) in js
                else:
                    assert int(ro[i, 2 * k + 1]) == -1
        seen += b.batch_size
    assert seen == 36
    # ranks outside [1, max_rank] are ignored
    ro = build_rank_offset(torch.tensor([5, 5, 5]), torch.tensor([1, 9, 2]), 3)
    assert ro[1].tolist() == [-1] * 7 and ro[0].tolist() == [1, 1, 0, 2, 2, -1, -1]
    # only cmatch 222 / 223 ads take part when cmatch is given (data_feed.cc:3557-3560)
    ro = build_rank_offset(torch.tensor([5, 5, 5]), torch.tensor([1, 2, 3]), 3, cmatch=torch.tensor([222, 100, 223]))
    assert ro[1].tolist() == [-1] * 7 and ro[0].tolist() == [1, 1, 0, -1, -1, 3, 2] and ro[2, 0] == 3


def test_merge_by_ins_id(tmp_path):
    """Two half-instances with the same ins_id (disjoint slots) become one; groups of the wrong
    size and groups that fill one sparse slot twice are dropped."""
    import torch
    from paddlebox_b200.data_generator import format_instance
    from paddlebox_b200.dataset import DatasetFactory, dense_slot, sparse_slot
    lines = [
        format_instance([[1.0], [11], [0]], ins_id="a"),         # a: slot s0 here (a zero feasign = empty slot) ...
        format_instance([[0.0], [0], [21, 22]], ins_id="a"),     # ... slot s1 there -> merged
        format_instance([[1.0], [31], [32]], ins_id="b"),        # b: single record -> group size 1 != 2: dropped
        format_instance([[0.0], [41], [0]], ins_id="c"),
        format_instance([[1.0], [42], [43]], ins_id="c"),        # c: s0 filled twice -> conflict: dropped
    ]
    path = tmp_path / "part-0"
    path.write_text("\n".join(lines) + "\n")
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var([dense_slot("label", 1), sparse_slot("s0"), sparse_slot("s1")])
    ds.set_batch_size(4)
    ds.set_thread(1)
    ds.set_parse_ins_id(True)
    ds.set_merge_by_lineid(2)
    ds.set_filelist([str(path)])
    ds.load_into_memory()
    assert ds.get_memory_data_size() == 1
    ds.prepare_train(1)
    (b,) = list(ds.batches(device=torch.device("cpu"), with_ins_id=True))
    assert b.batch_size == 1 and b.ins_ids == ["a"]
    assert sorted(b.keys.tolist()) == [11, 21, 22]
    assert float(b.dense[0, 0]) == 1.0           # the first non-zero label copy wins


def test_compact_arena_pack_matches_record_chase(tmp_path):
    """Record-major packing from the compact arena equals packing from the heap records, also
    after a slot shuffle (arena rebuilt) and with the arena switched off."""
    import numpy as np
    from paddlebox_b200.dataset import DatasetFactory
    from tests.util_data import slot_vars, write_files

    def packed(ds):
        nat, out = ds._native, []
        for b in range(nat.num_batches(0)):
            B, n = nat.batch_shape(0, b)
            keys = np.zeros(n, dtype=np.uint64)
            off = np.zeros(B * (nat.n_sparse() + 1), dtype=np.int32)
            dense = np.zeros(B * nat.dense_dim(), dtype=np.float32)
            nat.pack_batch(0, b, keys.ctypes.data, off.ctypes.data, dense.ctypes.data, False)
            out.append((keys, off, dense))
        return out

    def make(mode):
        ds = DatasetFactory().create_dataset("InMemoryDataset")
        ds.set_use_var(slot_vars(5, 3))
        ds.set_batch_size(16)
        ds.set_thread(1)
        ds.set_merge_thread_num(1)
        ds.set_compact_batches(mode)
        ds.set_filelist(write_files(str(tmp_path), n_files=1, lines_per_file=70, n_sparse=5, n_dense=3, key_space=50))
        ds.load_into_memory()
        ds.prepare_train(1)
        return ds
    a, b = make(2), make(0)
    for (k1, o1, d1), (k2, o2, d2) in zip(packed(a), packed(b)):
        assert np.array_equal(k1, k2) and np.array_equal(o1, o2) and np.array_equal(d1, d2)
    assert len(packed(a)) == 5                       # 70 instances -> 4 x 16 + 6
    a.slots_shuffle(["slot1", "slot3"])
    b.slots_shuffle(["slot1", "slot3"])
    for (k1, o1, d1), (k2, o2, d2) in zip(packed(a), packed(b)):
        assert np.array_equal(k1, k2) and np.array_equal(o1, o2) and np.array_equal(d1, d2)


def test_fast_float_parser_matches_float32(tmp_path):
    """The hand-rolled decimal reader gives the float32 nearest to the text for generator-style
    values, and exotic spellings still go through strtof."""
    import numpy as np
    import torch
    from paddlebox_b200.dataset import DatasetFactory, dense_slot, sparse_slot
    rng = np.random.default_rng(0)
    vals = [f"{v:.{int(d)}f}" for v, d in zip(rng.normal(0, 100, 300), rng.integers(0, 9, 300))]
    vals += ["1e-3", "-2.5E+2", "+7", "0.000001", "123456789012345678", "3.", "-0.0", "1e10"]
    n = len(vals)
    path = tmp_path / "part-0"
    path.write_text(f"{n} " + " ".join(vals) + " 1 5\n")
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var([dense_slot("f", n), sparse_slot("s")])
    ds.set_batch_size(1)
    ds.set_thread(1)
    ds.set_filelist([str(path)])
    ds.load_into_memory()
    assert ds.get_memory_data_size() == 1
    ds.prepare_train(1)
    (b,) = list(ds.batches(device=torch.device("cpu")))
    want = np.array([np.float32(float(v)) for v in vals], dtype=np.float32)
    got = b.dense[0].numpy()
    assert np.array_equal(got, want), np.flatnonzero(got != want)


def test_pass_file_roundtrip_is_batch_identical(tmp_path):
    """save_pass_file / load_pass_file: same batches (both layouts), same unique keys, no records needed."""
    import time
    from paddlebox_b200.dataset import DatasetFactory
    files = write_files(str(tmp_path), n_files=2, lines_per_file=300, n_sparse=5, n_dense=3, key_space=200)

    def make():
        ds = DatasetFactory().create_dataset("InMemoryDataset")
        ds.set_use_var(slot_vars(5, 3))
        ds.set_batch_size(32)
        ds.set_thread(2)
        return ds

    a = make()
    a.set_filelist(files)
    a.load_into_memory()
    a.prepare_train(2, 0, shuffle=True)
    want = {w: [(b.keys.clone(), b.lod.clone(), b.dense.clone()) for b in a.batches(worker=w)] for w in (0, 1)}
    a.save_pass_file(str(tmp_path / "pass.arena"))
    b = make()
    b.load_pass_file(str(tmp_path / "pass.arena"))
    assert b.get_memory_data_size() == 600 and np.array_equal(np.sort(b.unique_keys()), np.sort(a.unique_keys()))
    b.prepare_train(2, 0, shuffle=False)
    for w in (0, 1):
        got = list(b.batches(worker=w))
        assert len(got) == len(want[w])
        for g, (k, l, d) in zip(got, want[w]):
            assert torch.equal(g.keys, k) and torch.equal(g.lod, l) and torch.equal(g.dense, d)
    # record-major staging buffers (the pipelined feed's input) agree as well
    ha = [(h.n_keys, h.batch_size) for h in a.host_batches(worker=0)]
    hb = [(h.n_keys, h.batch_size) for h in b.host_batches(worker=0)]
    assert ha == hb and len(ha) > 0
    # a shuffle permutes whole batches: same multiset of batches
    b.prepare_train(2, 0, shuffle=True)
    sums = sorted(int(x.keys.sum()) for w in (0, 1) for x in b.batches(worker=w))
    assert sums == sorted(int(k.sum()) for w in (0, 1) for k, _, _ in want[w])
    # record-level operations are refused, a wrong slot set too
    with pytest.raises(N.NativeError):
        b._native.slots_shuffle([0], 1)
    c = DatasetFactory().create_dataset("InMemoryDataset")
    c.set_use_var(slot_vars(4, 3))
    with pytest.raises(N.NativeError, match="sparse slots"):
        c.load_pass_file(str(tmp_path / "pass.arena"))
    # loading text again replaces the mapped pass
    b.set_filelist(files[:1])
    b.load_into_memory()
    assert b.get_memory_data_size() == 300
    b.release_memory()


def test_native_util_selftest(tmp_path):
    """C++-level test of the blocking channel and the thread pool (the reference tests these with gtests)."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = tmp_path / "util_selftest"
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-pthread", "-o", str(exe),
                           os.path.join(root, "csrc", "host", "tests", "util_selftest.cc")])
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "util selftest ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.skipif(not os.environ.get("PBX_RUN_SANITIZERS"), reason="set PBX_RUN_SANITIZERS=1 (two sanitizer builds, ~2 min)")
def test_host_runtime_under_sanitizers(tmp_path):
    """ASan + UBSan and TSan builds of csrc/host driven through the C ABI (benchmarks/sanitize_host.sh); the
    committed logs of the same run are profiles/sanitizer_host_{asan,tsan}.log."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([os.path.join(root, "benchmarks", "sanitize_host.sh"), str(tmp_path)], capture_output=True,
                       text=True, timeout=1200)
    assert r.returncode == 0, r.stdout + r.stderr
    for mode in ("asan", "tsan"):
        log = (tmp_path / f"sanitizer_host_{mode}.log").read_text()
        assert "exit code 0" in log and "WARNING" not in log and "ERROR" not in log, log


def test_shrink_compacts_in_place_and_keeps_survivors_intact():
    """Half of the rows fall under the threshold: exactly they disappear, the survivors keep their rows
    (aged by one day), lookups of the evicted keys miss."""
    import torch
    from paddlebox_b200.host import HostTable
    t = HostTable(embedx_dim=8)
    keys = np.arange(1, 60001, dtype=np.uint64)
    t.fetch_rows(keys, create=True)
    push = torch.zeros(keys.size, t.layout.push_stride)
    push[:, 1] = torch.from_numpy((keys % 10).astype(np.float32))
    push[:, 3:] = 0.5
    t.push(keys, push)
    before = t.fetch_rows(keys, create=False).copy()
    cfg = t.config
    score = cfg.nonclk_coeff * (before[:, 0] - before[:, 1]) + cfg.clk_coeff * before[:, 1]
    kept = t.shrink(1.0, 0.45, 30.0)
    want = keys[score >= 0.45]
    assert 0 < want.size < keys.size and kept == want.size and t.size() == want.size
    assert np.array_equal(np.sort(t.keys()), want)
    b = before[score >= 0.45]
    b[:, t.layout.off_tail + 2] += 1
    assert np.allclose(t.fetch_rows(want, create=False), b)
    assert t.size() == want.size                         # create=False lookups of evicted keys add nothing
    gone = keys[score < 0.45][:100]
    assert np.all(t.pull(gone).numpy() == 0)


def test_reader_handles_block_boundaries_long_lines_and_missing_final_newline(tmp_path):
    """The block reader carries unfinished lines across 4 MB reads, grows for a line longer than its buffer,
    takes a last line without a newline, counts empty / damaged lines as bad, and keeps CRLF files readable."""
    vars_ = slot_vars(2, 1)
    # (label, dense 1, slot0, slot1); slot1 of the long line carries 600 000 ids: one line of ~5 MB > the 4 MB buffer
    long_ids = " ".join(str(1000 + i) for i in range(600_000))
    lines = [f"1 1 1 0.5 1 {10 + i} 2 {20 + i} {30 + i}" for i in range(200_000)]      # ~6 MB: crosses a block boundary
    lines.insert(1234, "")                                                      # empty line
    lines.insert(5000, "1 1 1 0.5 1")                                           # cut line
    lines.insert(7777, f"1 0 1 0.25 1 7 600000 {long_ids}")
    p = tmp_path / "big.txt"
    p.write_text("\r\n".join(lines))                                            # CRLF, no newline at the end
    d = _ds(vars_)
    d.set_filelist([str(p)])
    d.load_into_memory()
    assert d.num_records() == 200_001
    st = d.stats()
    assert "lines=200003" in st and "bad_lines=2" in st, st
    uk = set(d.unique_keys().tolist())
    assert {0, 7, 10, 20, 30, 1000, 1000 + 599_999, 10 + 199_999, 30 + 199_999} <= uk
    assert uk == {0, 7} | set(range(10, 601_000))                              # 10..200029 and the long line's 1000..600999
