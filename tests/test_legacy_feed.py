"""Record-based legacy feeds (legacy_feed.py): wire format, text parsing, merge by instance id, page views,
global shuffle over two ranks, bridge to SlotBatch."""
import struct

import numpy as np
import pytest
import torch

from paddlebox_b200 import legacy_feed as LF


def _rec(ins, u, f=(), sid=0, rank=0, cmatch=0):
    return LF.Record(np.array([s for s, _ in u], np.uint64), np.array([sl for _, sl in u], np.uint16),
                     np.array([v for v, _ in f], np.float32), np.array([sl for _, sl in f], np.uint16), ins,
                     search_id=sid, rank=rank, cmatch=cmatch)


def test_record_wire_format_is_the_reference_layout():
    r = _rec("abc", [(0x1122334455667788, 3), (7, 0)], [(1.5, 5)])
    raw = LF.serialize_record(r)
    # u64 count | (u64 sign, f32 view, u16 slot) x 2 | u64 count | item | u64 len | bytes   (data_feed.h:1536-1583)
    assert len(raw) == 8 + 2 * 14 + 8 + 14 + 8 + 3
    assert struct.unpack_from("<Q", raw, 0)[0] == 2
    sign, fview, slot = struct.unpack_from("<QfH", raw, 8)
    assert sign == 0x1122334455667788 and slot == 3
    assert struct.pack("<f", fview) == struct.pack("<I", 0x55667788)          # float view = low 4 bytes of the union
    off = 8 + 28
    assert struct.unpack_from("<Q", raw, off)[0] == 1
    fsign, fv, fslot = struct.unpack_from("<QfH", raw, off + 8)
    assert fv == 1.5 and fslot == 5 and fsign == struct.unpack("<I", struct.pack("<f", 1.5))[0]
    assert raw[-3:] == b"abc"
    back = LF.deserialize_records(raw + LF.serialize_record(_rec("", [(9, 1)])))
    assert len(back) == 2 and back[0].ins_id == "abc" and back[0].u64_signs.tolist() == [0x1122334455667788, 7]
    assert back[0].f32_vals.tolist() == [1.5] and back[1].u64_slots.tolist() == [1]
    with pytest.raises(ValueError):
        LF.deserialize_records(raw[:-1])


def test_text_parse_merge_and_pv(tmp_path):
    types = ["uint64", "uint64", "float"]
    r = LF.parse_multislot_line("1 id7 2 11 0 1 22 2 0.5 0.0", types, parse_ins_id=True)
    assert r.ins_id == "id7" and r.u64_signs.tolist() == [11, 22] and r.u64_slots.tolist() == [0, 1]    # zero dropped
    assert r.f32_vals.tolist() == [0.5] and r.f32_slots.tolist() == [2]
    with pytest.raises(ValueError):
        LF.parse_multislot_line("0 1 5 1 0.1", types)
    f = tmp_path / "part-0"
    f.write_text("1 a 1 11 1 21 1 0.5\n1 a 1 12 1 22 1 0.7\n1 b 1 13 1 0 1 0.1\n1 c 1 14 1 24 1 0.2\n1 c 1 15 1 25 1 0.3\n")
    ds = LF.MultiSlotDataset(types, batch_size=2, used=[True, False, True])
    ds.set_filelist([str(f)])
    ds.set_merge_by_lineid(2)
    assert ds.load_into_memory() == 5
    dropped = ds.merge_by_lineid()
    # a / c: two records each, but both carry slot 0 -> conflicting groups are dropped; b has the wrong group size
    assert dropped == 5 and ds.get_memory_data_size() == 0
    ds2 = LF.MultiSlotDataset(types, batch_size=2)
    g = tmp_path / "part-1"
    g.write_text("1 a 1 11 1 0 1 0.5\n1 a 1 0 1 22 1 0.7\n1 b 1 13 1 23 1 0.1\n")
    ds2.set_filelist([str(g)])
    ds2.set_merge_by_lineid(0)
    ds2.load_into_memory()
    assert ds2.merge_by_lineid() == 0 and ds2.get_memory_data_size() == 2
    a = [r for r in ds2.records if r.ins_id == "a"][0]
    assert sorted(a.u64_signs.tolist()) == [11, 22] and a.f32_vals.tolist() == [0.5]      # first float copy wins
    # page views
    ds2.records = [_rec("x", [(1, 0)], sid=5, rank=2, cmatch=222), _rec("y", [(2, 0)], sid=5, rank=1, cmatch=222),
                   _rec("z", [(3, 0)], sid=9, rank=1, cmatch=223)]
    assert ds2.preprocess_instance() == 2
    recs, ro = next(ds2.pv_batches(2))
    assert [r.ins_id for r in recs] == ["y", "x", "z"] and ro.shape == (3, 7) and ro[0, 0] == 1 and ro[1, 0] == 2
    ds2.postprocess_instance()
    assert [r.ins_id for r in ds2.records] == ["y", "x", "z"]
    # streaming feed
    feed = LF.MultiSlotDataFeed(types, batch_size=2, parse_ins_id=True)
    bs = list(feed.batches([str(g)]))
    assert [len(b) for b in bs] == [2, 1]


def test_records_to_slot_batch_trains_on_cpu():
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer
    recs = [_rec(str(i), [(100 + i, 0), (200 + i % 3, 1), (300 + i, 1)], [(float(i % 2), 2), (0.25 * i, 3)]) for i in range(8)]
    b = LF.records_to_slot_batch(recs, sparse_slots=[0, 1], dense_slots=[2, 3])
    assert b.batch_size == 8 and b.num_slots == 2 and b.lod.tolist()[:9] == list(range(9)) and b.lod[-1] == 24
    assert b.keys[:8].tolist() == list(range(100, 108)) and b.dense.shape == (8, 2) and b.dense[3].tolist() == [1.0, 0.75]
    BoxWrapper.reset_instance()
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    try:
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
        box.begin_feed_pass("20240101"); box.feed_pass(b.keys); box.end_feed_pass(); box.begin_pass()
        torch.manual_seed(0)
        tr = CTRTrainer(CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(8,)), box)
        losses = [float(tr.train_step(b).detach()) for _ in range(80)]
        box.end_pass()
        assert all(np.isfinite(losses)) and min(losses[-10:]) < losses[0]
    finally:
        BoxWrapper.reset_instance()
