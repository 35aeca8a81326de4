"""Property tests (hypothesis) of host-side invariants: wire formats round-trip, parsers agree with the
generator, the two gradient merges agree, dedup is a permutation inverse, batch dealing loses nothing,
the native sort / partition agree with numpy.  CPU only, small examples, fixed budget."""
import numpy as np
import pytest
import torch

hyp = pytest.importorskip("hypothesis")
from hypothesis import HealthCheck, given, settings  # noqa: E402
from hypothesis import strategies as st  # noqa: E402

FAST = settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.too_slow])

u64s = st.integers(min_value=1, max_value=(1 << 64) - 1)
floats = st.floats(min_value=-1e6, max_value=1e6, allow_nan=False, width=32).filter(lambda v: abs(v) >= 1e-3)


# ------------------------------------------------------------------------------- Record wire format
@st.composite
def records(draw):
    from paddlebox_b200.legacy_feed import Record
    nu = draw(st.integers(0, 12))
    nf = draw(st.integers(0, 6))
    return Record(np.array(draw(st.lists(u64s, min_size=nu, max_size=nu)), np.uint64),
                  np.array(draw(st.lists(st.integers(0, 65535), min_size=nu, max_size=nu)), np.uint16),
                  np.array(draw(st.lists(floats, min_size=nf, max_size=nf)), np.float32),
                  np.array(draw(st.lists(st.integers(0, 65535), min_size=nf, max_size=nf)), np.uint16),
                  draw(st.text(alphabet="abcdefXYZ0123456789_-", max_size=24)))


@FAST
@given(st.lists(records(), max_size=6))
def test_record_archive_round_trips(recs):
    from paddlebox_b200.legacy_feed import deserialize_records, serialize_records
    buf = serialize_records(recs)
    back = deserialize_records(buf)
    assert len(back) == len(recs)
    for a, b in zip(recs, back):
        assert np.array_equal(a.u64_signs, b.u64_signs) and np.array_equal(a.u64_slots, b.u64_slots)
        assert np.array_equal(a.f32_vals, b.f32_vals) and np.array_equal(a.f32_slots, b.f32_slots)
        assert a.ins_id == b.ins_id
    # any strict prefix that cuts a record is refused, never mis-read
    if len(buf) > 9:
        cut = buf[:len(buf) - 3]
        try:
            got = deserialize_records(cut)
            assert len(got) < len(recs) or len(recs) == 0
        except (ValueError, Exception):
            pass


# ------------------------------------------------------------------------------- text format
@st.composite
def instance(draw):
    types = draw(st.lists(st.sampled_from(["uint64", "float"]), min_size=1, max_size=6))
    slots = []
    for ty in types:
        n = draw(st.integers(1, 4))
        slots.append(draw(st.lists(u64s if ty == "uint64" else floats, min_size=n, max_size=n)))
    return types, slots


@FAST
@given(instance())
def test_multislot_text_line_parses_back_to_what_was_written(inst):
    from paddlebox_b200.legacy_feed import parse_multislot_line
    types, slots = inst
    line = " ".join(f"{len(v)} " + " ".join(str(x) if ty == "uint64" else repr(float(x)) for x in v)
                    for ty, v in zip(types, slots))
    r = parse_multislot_line("1 id7 " + line, types, parse_ins_id=True)
    assert r.ins_id == "id7"
    want_u = [(x, i) for i, (ty, v) in enumerate(zip(types, slots)) if ty == "uint64" for x in v]
    want_f = [(np.float32(x), i) for i, (ty, v) in enumerate(zip(types, slots)) if ty == "float" for x in v]
    assert [int(x) for x in r.u64_signs] == [x for x, _ in want_u] and list(r.u64_slots) == [i for _, i in want_u]
    assert np.array_equal(r.f32_vals, np.array([x for x, _ in want_f], np.float32))
    assert list(r.f32_slots) == [i for _, i in want_f]
    # a line cut in the middle of a slot is an error, not a shorter record
    toks = line.split()
    if len(toks) > 2:
        with pytest.raises(ValueError):
            parse_multislot_line(" ".join(toks[:-1]), types)


@FAST
@given(st.lists(st.lists(u64s, min_size=1, max_size=4), min_size=1, max_size=5),
       st.lists(floats, min_size=2, max_size=2))
def test_native_slot_parser_agrees_with_the_generator(sparse, dense):
    """format_instance -> native text parser (the SlotRecord path) -> the same values, slot by slot."""
    import ctypes as C
    from paddlebox_b200.data_generator import format_instance
    from paddlebox_b200.dataset import DatasetFactory, dense_slot, sparse_slot
    import os
    import tempfile
    line = format_instance([[1.0], [round(float(d), 4) for d in dense]] + sparse)
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "f")
        with open(path, "w") as f:
            f.write(line + "\n")
        ds = DatasetFactory().create_dataset("InMemoryDataset")
        ds.set_use_var([dense_slot("label", 1), dense_slot("d", 2)] + [sparse_slot(f"s{i}") for i in range(len(sparse))])
        ds.set_batch_size(1)
        ds.set_filelist([path])
        ds.load_into_memory()
        assert ds.get_memory_data_size() == 1
        ds.prepare_train(1)
        hb = next(iter(ds.host_batches()))
        keys = hb.keys[:hb.n_keys].numpy().view(np.uint64)
        assert sorted(int(k) for k in keys) == sorted(x for v in sparse for x in v)
        uk = set(int(k) for k in ds.unique_keys())
        assert uk == {0} | {x for v in sparse for x in v}
        hb.release()
        ds.release_memory()


# ------------------------------------------------------------------------------- gradient merges / dedup
@FAST
@given(st.integers(1, 60), st.integers(1, 12), st.integers(1, 6), st.integers(0, 2 ** 31 - 1))
def test_sorted_and_atomic_merges_agree(n, n_unique, dim, seed):
    from paddlebox_b200.ops import reference as R
    g = torch.Generator().manual_seed(seed)
    restore = torch.randint(0, n_unique, (n,), generator=g, dtype=torch.int32)
    grads = torch.randn(n, 3 + dim, generator=g)
    grads[:, 0] = 1.0                                   # the show column of an occurrence is 1
    key2slot = torch.randint(0, 5, (n,), generator=g, dtype=torch.int32)
    a = R.merge_push(restore, n_unique, grads, key2slot, 2, -4.0)
    b = R.merge_push_sorted(restore, n_unique, grads, key2slot, 2, -4.0)
    assert a.shape == b.shape
    touched = torch.zeros(n_unique, dtype=torch.bool)
    touched[restore.long()] = True
    assert torch.allclose(a[touched][:, 1:], b[touched][:, 1:], atol=1e-4, rtol=1e-4)
    assert torch.equal(b[~touched][:, 1:], torch.zeros_like(b[~touched][:, 1:]))
    # show column (not scaled: it is below cvm_offset) = number of occurrences of the key
    cnt = torch.bincount(restore.long(), minlength=n_unique).float()
    assert torch.equal(b[:, 1], cnt)


@FAST
@given(st.lists(st.integers(0, 30), min_size=1, max_size=80))
def test_dedup_restore_is_an_inverse(keys):
    from paddlebox_b200.ops import reference as R
    k = torch.tensor(keys, dtype=torch.int64)
    uniq, restore = R.dedup_keys_and_fillidx(k)[:2]
    assert torch.equal(uniq[restore.long()], k)
    assert uniq.numel() == len(set(keys))


# ------------------------------------------------------------------------------- native helpers vs numpy
@FAST
@given(st.lists(st.integers(0, (1 << 64) - 1), max_size=300), st.integers(1, 9))
def test_native_sort_unique_and_partition_agree_with_numpy(keys, world):
    from paddlebox_b200.box_wrapper import owner_of
    from paddlebox_b200.host import partition_by_owner, sort_unique_u64
    a = np.array(keys, np.uint64)
    u = sort_unique_u64(a.copy())
    assert np.array_equal(u, np.unique(a))
    parts = partition_by_owner(u, world)
    assert len(parts) == world and sum(p.size for p in parts) == u.size
    for r, p in enumerate(parts):
        if p.size:
            assert np.all(owner_of(p, world) == r)
    assert np.array_equal(np.sort(np.concatenate(parts)) if u.size else u, u)


@FAST
@given(st.integers(0, 500), st.integers(1, 64), st.integers(1, 8))
def test_batch_dealing_covers_every_instance_once(n_ins, batch, workers):
    """PrepareTrain's dealing (data_set.cc:2660-2750): every instance lands in exactly one batch, no batch is
    larger than the batch size, and every worker gets the same number of batches when there are enough."""
    import ctypes as C
    from paddlebox_b200 import _native as N
    lib = N.host()
    if not hasattr(lib, "pbh_compute_batch_ranges"):
        pytest.skip("batch range helper not exported")
    cap = n_ins + workers + 8
    starts = (C.c_int64 * cap)()
    lens = (C.c_int32 * cap)()
    n = lib.pbh_compute_batch_ranges(C.c_int64(n_ins), batch, workers, starts, lens, cap)
    assert n >= 0
    seen = np.zeros(n_ins, np.int32)
    per_worker_ok = True
    for i in range(n):
        # (a worker may get an empty batch when there are fewer instances than workers, and the first batch of
        # a folded tail round carries the remainder: both as in the reference's compute_batch_num)
        assert 0 <= lens[i] <= batch + 2 * workers
        seen[starts[i]:starts[i] + lens[i]] += 1
    assert np.all(seen == 1)
    assert n % workers == 0                                  # whole rounds: the same number of batches per worker


# ------------------------------------------------------------------------------- protobuf text format
_names = st.text(alphabet="abcdefghijklmnopqrstuvwxyz_", min_size=1, max_size=8).filter(lambda s: s not in ("true", "false", "inf", "nan"))
_scalars = st.one_of(st.integers(-10 ** 9, 10 ** 9), st.booleans(),
                     st.floats(min_value=-1e6, max_value=1e6, allow_nan=False).map(lambda v: round(v, 4)).filter(lambda v: v != int(v)),
                     st.text(alphabet="abcXYZ012 _-./:{}#", max_size=12))


def _messages(depth):
    if depth == 0:
        return st.dictionaries(_names, _scalars, max_size=4)
    return st.dictionaries(_names, st.one_of(_scalars, _messages(depth - 1), st.lists(_scalars, min_size=2, max_size=3)), max_size=4)


def _to_text(msg, indent=0):
    pad = " " * indent
    out = []
    for k, v in msg.items():
        items = v if isinstance(v, list) else [v]
        for it in items:
            if isinstance(it, dict):
                out.append(f"{pad}{k} {{\n{_to_text(it, indent + 2)}{pad}}}\n")
            elif isinstance(it, bool):
                out.append(f"{pad}{k}: {'true' if it else 'false'}\n")
            elif isinstance(it, str):
                out.append(pad + k + ': "' + it.replace("\\", "\\\\").replace('"', '\\"') + '"\n')
            else:
                out.append(f"{pad}{k}: {it!r}\n")
    return "".join(out)


@FAST
@given(_messages(2))
def test_protobuf_text_parser_round_trips_nested_messages(msg):
    from paddlebox_b200.fleet_desc import parse_text_format
    text = _to_text(msg)
    got = parse_text_format(text)

    def same(a, b):
        if isinstance(a, dict):
            return isinstance(b, dict) and a.keys() == b.keys() and all(same(a[k], b[k]) for k in a)
        if isinstance(a, list):
            return isinstance(b, list) and len(a) == len(b) and all(same(x, y) for x, y in zip(a, b))
        if isinstance(a, float):
            return abs(a - b) < 1e-9
        return a == b and type(a) is type(b)
    assert same(msg, got), (msg, got, text)
    from paddlebox_b200.fleet_desc import format_text
    again = parse_text_format(format_text(msg))                   # the framework's own writer reads back the same
    assert same(msg, again), (msg, again, format_text(msg))


@FAST
@given(st.sampled_from([1, 2, 3, 5, 6, 10, 15, 30, 60]), st.integers(1, 6), st.integers(0, 23), st.integers(0, 23), st.booleans())
def test_online_pass_intervals_cover_the_allowed_hours_in_order(split_interval, per_pass, h0, h1, hourly):
    from paddlebox_b200.fleet_util import FleetUtil
    lo, hi = min(h0, h1), max(h0, h1)
    passes = FleetUtil().get_online_pass_interval("{20240101..20240102}", "{%d..%d}" % (lo, hi), split_interval, per_pass, hourly)
    flat = [s for p in passes for s in p]
    assert all(len(p) == per_pass for p in passes)
    want = []
    for i in range(24 * 60 // split_interval):
        h, m = divmod(i * split_interval, 60)
        if lo <= h <= hi:
            want.append("%02d" % h if hourly else "%02d%02d" % (h, m))
    assert flat == want[:len(flat)] and len(want) - len(flat) < per_pass + max(0, len(want) - (24 * 60 // split_interval // per_pass) * per_pass)
