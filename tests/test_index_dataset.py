"""Tree index (TDM): file format, queries, layer-wise sampler, and the sampling of a loaded dataset."""
import numpy as np
import pytest

from paddlebox_b200.index_dataset import IndexNode, IndexSampler, IndexWrapper, TreeIndex, save_tree_index


def _write_tree(path, n_leaf=8, branch=2, missing=()):
    """A complete tree over ``n_leaf`` items: leaf i has id 100 + i; inner node with code c has id 1000 + c."""
    height = 1
    while branch ** (height - 1) < n_leaf:
        height += 1
    first_leaf = (branch ** (height - 1) - 1) // (branch - 1)
    nodes = {c: IndexNode(1000 + c, False, 0.5) for c in range(first_leaf)}
    for i in range(n_leaf):
        if i not in missing:
            nodes[first_leaf + i] = IndexNode(100 + i, True, 0.25, f"item{i}")
    save_tree_index(str(path), nodes, height, branch)
    return height, first_leaf


@pytest.fixture(autouse=True)
def _fresh_registry():
    IndexWrapper().clear_tree()
    yield
    IndexWrapper().clear_tree()


def test_tree_file_round_trip_and_queries(tmp_path):
    height, first_leaf = _write_tree(tmp_path / "tree.pb")
    t = TreeIndex("t", str(tmp_path / "tree.pb"))
    assert (t.height(), t.branch(), t.total_node_nums()) == (4, 2, 15) and t.emb_size() == 1007
    leafs = t.get_all_leafs()
    assert sorted(n.id() for n in leafs) == list(range(100, 108)) and all(n.is_leaf() for n in leafs)
    assert leafs[0].item_name().startswith("item") and abs(leafs[0].probability() - 0.25) < 1e-7
    assert t.get_layer_codes(0) == [0] and t.get_layer_codes(2) == [3, 4, 5, 6] and t.get_layer_codes(3) == list(range(7, 15))
    assert t.get_travel_codes(103, 0) == [10, 4, 1, 0] and t.get_travel_codes(103, 2) == [10, 4]
    with pytest.raises(ValueError):
        t.get_travel_codes(999)
    assert t.get_ancestor_codes([100, 107, 999], 1) == [1, 2, 15]                   # unknown id -> one past the largest code
    assert t.get_ancestor_codes([105], 3) == [12]
    assert t.get_children_codes(1, 3) == [7, 8, 9, 10] and t.get_children_codes(2, 2) == [5, 6] and t.get_children_codes(0, 1) == [1, 2]
    nodes = t.get_nodes([0, 7, 99])
    assert [n.id() for n in nodes] == [1000, 100, 0] and not nodes[2].is_leaf()      # a code without a node -> the empty node
    assert t.get_travel_path(10, 1) == [10, 4] and t.get_pi_relation([100, 104], 1) == {100: 1, 104: 2}
    tdm = t.to_tdm_tree()
    assert tdm.branch == 2


def test_sparse_and_ternary_trees(tmp_path):
    _write_tree(tmp_path / "gap.pb", n_leaf=8, missing=(2, 5))
    t = TreeIndex("gap", str(tmp_path / "gap.pb"))
    assert t.total_node_nums() == 13 and t.get_layer_codes(3) == [7, 8, 10, 11, 13, 14] and t.get_children_codes(1, 3) == [7, 8, 10]
    _write_tree(tmp_path / "tri.pb", n_leaf=9, branch=3)
    t3 = TreeIndex("tri", str(tmp_path / "tri.pb"))
    assert (t3.height(), t3.branch(), t3.total_node_nums()) == (3, 3, 13)
    assert t3.get_layer_codes(1) == [1, 2, 3] and t3.get_layer_codes(2) == list(range(4, 13))
    assert t3.get_travel_codes(104, 0) == [8, 2, 0] and t3.get_children_codes(2, 2) == [7, 8, 9]
    (tmp_path / "bad.pb").write_bytes(b"\x40\x00\x00\x00abc")
    with pytest.raises(ValueError):
        TreeIndex("bad", str(tmp_path / "bad.pb"))
    with pytest.raises(ValueError):
        IndexWrapper().get_tree_index("nosuch")


def test_layerwise_sampler_rows(tmp_path):
    _write_tree(tmp_path / "tree.pb")
    t = TreeIndex("t", str(tmp_path / "tree.pb"))
    with pytest.raises(ValueError):
        t.layerwise_sample([[100]], [101])
    with pytest.raises(ValueError):
        IndexSampler("by_layerwise", "t").init_layerwise_conf([1], start_sample_layer=4)
    t.init_layerwise_sampler([1, 2, 3], start_sample_layer=1, seed=7)             # layers 1, 2, 3 -> 1, 2, 3 negatives
    rows = t.layerwise_sample([[100, 101], [106, 107]], [103, 105], with_hierarchy=False)
    per = (3 + 1) + (2 + 1) + (1 + 1)
    assert len(rows) == 2 * per and all(len(r) == 4 for r in rows)
    first = rows[:per]
    assert all(r[:2] == [100, 101] for r in first)
    # leaf layer first: positive = the item itself, then 3 negatives that are other leaves
    assert first[0][2:] == [103, 1] and all(r[3] == 0 and 100 <= r[2] <= 107 and r[2] != 103 for r in first[1:4])
    assert first[4][2:] == [1004, 1] and all(r[3] == 0 and r[2] in (1003, 1005, 1006) for r in first[5:7])
    assert first[7][2:] == [1001, 1] and first[8][2:] == [1002, 0]
    h = t.layerwise_sample([[100, 101], [106, 107]], [103, 105], with_hierarchy=True)
    assert h[0][:2] == [100, 101] and h[4][:2] == [1003, 1003] and h[7][:2] == [1001, 1001]   # users lifted to the row's layer
    assert h[per + 4][:2] == [1006, 1006] and h[per + 7][:2] == [1002, 1002]
    assert IndexSampler("by_beamsearch", "t").sample([[1]], [100]) == []
    # negatives are uniform over the other nodes of the layer
    s = IndexSampler("by_layerwise", "t")
    s.init_layerwise_conf([0, 0, 4], 1, seed=1)
    codes, labels = s.sample_codes(np.full(3000, 103))
    neg = codes[:, 1:5].reshape(-1)
    assert 10 not in neg and set(neg.tolist()) == set(range(7, 15)) - {10}
    assert np.bincount(neg - 7, minlength=8)[[0, 1, 2, 4, 5, 6, 7]].min() > 3000 * 4 / 7 * 0.85
    assert labels[:, 0].all() and not labels[:, 1:5].any()


def test_dataset_tdm_sample_expands_records(tmp_path):
    from paddlebox_b200.dataset import DatasetFactory, dense_slot, sparse_slot
    _write_tree(tmp_path / "tree.pb")
    data = tmp_path / "part-0"
    # slots: user (sparse), item (sparse), label (dense float)
    lines = [f"2 {900 + i} {901 + i} 1 {100 + i % 8} 1 1.0\n" for i in range(20)]
    data.write_text("".join(lines))
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var([sparse_slot("user"), sparse_slot("item"), dense_slot("label", 1)])
    ds.set_batch_size(1000)
    ds.set_thread(1)
    ds.set_filelist([str(data)])
    ds.load_into_memory()
    keys_before = set(ds.unique_keys().tolist())
    n = ds.tdm_sample("t", str(tmp_path / "tree.pb"), [1, 1, 2], start_sample_layer=1, seed=3, id_slot="item")
    per = (2 + 1) + (1 + 1) + (1 + 1)
    assert n == 20 * per == ds.get_memory_data_size()
    keys_after = set(ds.unique_keys().tolist())
    assert keys_before <= keys_after and {1001, 1002} <= keys_after                # sampled node ids are pass keys now
    ds._native.prepare_train(1, False, 0, 0)
    vals, has = ds._native.slot_first_values(1)
    assert has.all() and set(vals.tolist()) <= set(range(100, 108)) | set(range(1001, 1007))
    # every source record: 3 positives (leaf, layer 2, layer 1) with label 1, 4 negatives with label 0
    assert int(np.isin(vals, list(range(1001, 1003))).sum()) == 20 * 2             # layer 1: one positive + one negative each
    import torch
    ds.prepare_train(1, shuffle=False) if hasattr(ds, "prepare_train") else None
    (b,) = list(ds.batches(worker=0, device=torch.device("cpu")))
    B = b.batch_size
    item = b.keys[b.lod[B:2 * B].long()]                                           # slot 1, one id per instance
    label = b.dense[:, 0]
    assert B == 20 * per and int(label.sum()) == 20 * 3                            # 3 positives per source record
    pos_items = item[label == 1].tolist()
    assert sum(100 <= v <= 107 for v in pos_items) == 20 and sum(v in (1001, 1002) for v in pos_items) == 20
    users, _ = ds._native.slot_first_values(0)
    assert sorted(set(users.tolist())) == [900 + i for i in range(20)]       # other slots ride along unchanged


def test_damaged_tree_files_are_refused_not_crashed_on(tmp_path):
    _write_tree(tmp_path / "tree.pb")
    good = (tmp_path / "tree.pb").read_bytes()
    rng = np.random.RandomState(0)
    outcomes = {"ok": 0, "refused": 0}
    for trial in range(300):
        buf = bytearray(good)
        if trial % 2:
            del buf[rng.randint(1, len(buf)):]                                 # truncated
        else:
            for _ in range(rng.randint(1, 4)):
                buf[rng.randint(0, len(buf))] ^= 1 << rng.randint(0, 8)        # flipped bits
        p = tmp_path / "damaged.pb"
        p.write_bytes(bytes(buf))
        IndexWrapper().clear_tree()
        try:
            t = TreeIndex("d", str(p))
            t.get_layer_codes(1), t.get_all_leafs(), t.get_ancestor_codes([100, 103], 1), t.get_children_codes(0, 2)
            outcomes["ok"] += 1
        except ValueError:
            outcomes["refused"] += 1
    assert outcomes["ok"] + outcomes["refused"] == 300 and outcomes["refused"] > 0
