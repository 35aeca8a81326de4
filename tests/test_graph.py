"""Graph engine: CSR build, host / device sampling and walks, skip-gram training on walk pairs."""
import numpy as np
import torch

from paddlebox_b200.graph import DeviceGraph, GraphDataGenerator, GraphGpuWrapper, GraphTable, walk_pairs


def _two_cliques(n=12):
    """Nodes 100..100+n-1 and 200..200+n-1, each group fully connected, one bridge 100 <-> 200."""
    src, dst = [], []
    for base in (100, 200):
        for a in range(n):
            for b in range(n):
                if a != b:
                    src.append(base + a)
                    dst.append(base + b)
    src += [100, 200]
    dst += [200, 100]
    return np.array(src, np.uint64), np.array(dst, np.uint64)


def test_csr_sampling_and_walks_host_and_device(tmp_path):
    src, dst = _two_cliques()
    g = GraphTable()
    g.add_edges(src, dst)
    g.build()
    ids, rp, col = g.csr()
    assert g.num_nodes() == 24 and g.num_edges() == src.size and np.all(np.diff(ids.astype(np.int64)) > 0)
    adj = {}
    for a, b in zip(src.tolist(), dst.tolist()):
        adj.setdefault(a, set()).add(b)
    for i, node in enumerate(ids.tolist()):
        assert set(ids[col[rp[i]:rp[i + 1]]].tolist()) == adj[node]
    # host sampling: without replacement, inside the neighbourhood, all of it when k >= degree
    nb, cnt = g.sample_neighbors(ids, 5, seed=3)
    for node, row, c in zip(ids.tolist(), nb, cnt):
        assert c == 5 and len(set(row.tolist())) == 5 and set(row.tolist()) <= adj[node]
    nb, cnt = g.sample_neighbors(np.array([105, 999], np.uint64), 20)
    assert cnt.tolist() == [11, 0] and set(nb[0, :11].tolist()) == adj[105] and not nb[1].any()
    # walks follow edges; an unknown start ends at once
    walks, lens = g.random_walk(np.array([100, 207, 999], np.uint64), 10, seed=1)
    assert lens.tolist() == [10, 10, 1] and walks[2, 1:].sum() == 0
    for w in walks[:2]:
        assert all(int(b) in adj[int(a)] for a, b in zip(w[:-1], w[1:]))
    # device twin (CPU tensors here; the same ops run on a GPU)
    dg = g.to_device("cpu")
    q = torch.tensor([100, 105, 999, 211], dtype=torch.int64)
    assert dg.degree(q).tolist() == [12, 11, 0, 11]
    gen = torch.Generator().manual_seed(0)
    nb, cnt = dg.sample_neighbors(q, 4, generator=gen)
    assert cnt.tolist() == [4, 4, 0, 4] and not nb[2].any()
    assert all(set(r.tolist()) <= adj[int(n)] for n, r in zip(q.tolist(), nb) if int(n) in adj)
    nb, cnt = dg.sample_neighbors(q, 16, generator=gen)
    assert cnt.tolist() == [12, 11, 0, 11] and set(nb[1, :11].tolist()) == adj[105] and not nb[1, 11:].any()
    walks, lens = dg.random_walk(q, 7, generator=gen)
    assert lens.tolist() == [7, 7, 1, 7]
    for w, l in zip(walks.tolist(), lens.tolist()):
        assert all(b in adj[a] for a, b in zip(w[:l - 1], w[1:l])) and not any(w[l:])
    c, x = walk_pairs(walks, lens, window=2)
    assert c.numel() == x.numel() == 3 * (2 * (7 * 2) - 2 * 3)          # 3 full walks, window 2
    # metapath over two edge types + the reference's wrapper names
    f1, f2 = tmp_path / "u2i.txt", tmp_path / "i2u.txt"
    f1.write_text("".join(f"{u}\t{1000 + u % 3}\n" for u in range(1, 10)))
    w = GraphGpuWrapper()
    w.set_up_types(["u2i", "i2u"])
    assert w.load_edge_file("u2i", str(f1)) == 9 and w.load_edge_file("i2u", str(f1), reverse=True) == 9
    dgraph = w.upload_batch("cpu")
    assert sorted(w.get_all_id("u2i").tolist()) == list(range(1, 10)) + [1000, 1001, 1002]
    walks, lens = dgraph.random_walk(torch.arange(1, 10), 5, metapath=(0, 1), generator=gen)
    assert lens.tolist() == [5] * 9
    assert torch.all(walks[:, 1::2] >= 1000) and torch.all(walks[:, 0::2] < 1000)
    assert torch.all((walks[:, 0] % 3) == (walks[:, 2] % 3))           # users sharing an item
    hw, hl = w.table.random_walk(np.arange(1, 10, dtype=np.uint64), 5, metapath=(0, 1))
    assert np.all(hl == 5) and np.all(hw[:, 1::2] >= 1000)
    nb, cnt = w.graph_neighbor_sample("i2u", np.array([1000], np.uint64), 2)
    assert cnt[0] == 2 and all(int(v) % 3 == 0 for v in nb[0])


def test_skipgram_trains_on_walk_pairs():
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.models.graph import SkipGram
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    BoxWrapper.reset_instance()
    box = BoxWrapper.set_instance(embedx_dim=8, device=torch.device("cpu"), rank=0, world_size=1)
    try:
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0, "mf_initial_range": 0.3,
                                           "mf_learning_rate": 0.2, "learning_rate": 0.05})
        src, dst = _two_cliques()
        g = GraphTable()
        g.add_edges(src, dst)
        g.build()
        gen = GraphDataGenerator(g, walk_len=6, window=2, batch_size=256, neg=2, walks_per_node=4, seed=1,
                                 device="cpu")
        box.begin_feed_pass("20240101")
        box.feed_pass(gen.all_keys())
        box.end_feed_pass()
        box.begin_pass()
        torch.manual_seed(0)
        tr = CTRTrainer(SkipGram(8), box, TrainerConfig(dense_lr=1e-2))
        losses = []
        for _ in range(12):
            for batch in gen.batches():
                losses.append(float(tr.train_step(batch).detach()))
        assert np.mean(losses[-10:]) < 0.8 * np.mean(losses[:10]), (losses[:3], losses[-3:])
        # nodes of one clique ended up closer to each other than to the other clique
        ids = torch.arange(100, 112).tolist() + torch.arange(200, 212).tolist()
        vals, restore = box.pull_sparse(torch.tensor(ids, dtype=torch.int64))
        emb = torch.nn.functional.normalize(vals[restore.long()][:, 3:11], dim=1)
        sim = emb @ emb.T
        inside = (sim[:12, :12].sum() - 12 + sim[12:, 12:].sum() - 12) / (2 * 12 * 11)
        across = sim[:12, 12:].mean()
        assert inside > across + 0.05, (float(inside), float(across))
        box.end_pass()
    finally:
        BoxWrapper.reset_instance()


def test_weighted_edges_node_features_and_partitions(tmp_path):
    """The rest of the graph engine's surface (graph_gpu_wrapper.h:50-99): weighted sampling, node files with
    feature slots, node weights, partitions for paging."""
    from paddlebox_b200.graph import GraphGpuWrapper, GraphTable
    g = GraphTable()
    # node 1 has 6 neighbours; neighbour 16 carries almost all the weight
    dst = np.array([11, 12, 13, 14, 15, 16], np.uint64)
    w = np.array([1, 1, 1, 1, 1, 500], np.float32)
    g.add_weighted_edges(np.full(6, 1, np.uint64), dst, w)
    g.add_weighted_edges(np.array([2, 2], np.uint64), np.array([11, 12], np.uint64), np.array([1.0, 3.0], np.float32))
    g.build(0)
    assert g.is_weighted(0)
    ids, rp, col = g.csr(0)
    ew = g.edge_weights(0)
    i1 = int(np.searchsorted(ids, 1))
    assert sorted(ew[rp[i1]:rp[i1 + 1]].tolist()) == [1, 1, 1, 1, 1, 500]
    hits = 0
    for seed in range(200):
        out, cnt = g.sample_neighbors_weighted(np.array([1], np.uint64), 2, 0, seed)
        assert cnt[0] == 2 and len(set(out[0].tolist())) == 2 and set(out[0].tolist()) <= set(dst.tolist())
        hits += 16 in out[0].tolist()
    assert hits > 190                                   # the heavy neighbour is (nearly) always drawn
    out, cnt = g.sample_neighbors_weighted(np.array([2, 99], np.uint64), 4, 0, 1)
    assert cnt.tolist() == [2, 0] and sorted(out[0][:2].tolist()) == [11, 12]
    # node file with feature slots
    nf = tmp_path / "nodes.txt"
    nf.write_text("user\t1\t101:7001\t102:8001 8002\nuser\t2\t101:7002\nitem\t11\t201:9001\nuser\t3\n")
    assert g.load_node_file(str(nf), 0, "user", " ") == 3
    assert g.load_node_file(str(nf), 1, "item", " ") == 1
    assert g.typed_nodes(0).tolist() == [1, 2, 3] and g.typed_nodes(1).tolist() == [11]
    f = g.node_features(np.array([2, 1, 3, 77], np.uint64), [101, 102], 0)
    assert f.tolist() == [[7002, 0], [7001, 8001], [0, 0], [0, 0]]
    assert g.all_feature_ids(0).tolist() == [7001, 7002, 8001, 8002]
    wf = tmp_path / "w.txt"
    wf.write_text("1\t2.5\n3\t0.5\n42\t9\n")
    assert g.load_node_weight(str(wf), 0) == 2
    assert g.node_weights(np.array([1, 2, 3, 42], np.uint64), 0).tolist() == [2.5, 1.0, 0.5, 0.0]
    # partitions: a chain graph cut into pieces by a byte budget
    h = GraphTable()
    n = 1000
    h.add_edges(np.arange(n, dtype=np.uint64), (np.arange(n, dtype=np.uint64) + 1) % n)
    h.build(0)
    parts = h.make_partitions(0, 32 * 100, 1)           # 32 bytes per node with one edge -> 100 nodes per partition
    assert parts == 10 and h.partition_num(0) == 10
    seen = []
    while True:
        p = h.next_partition(0)
        if p < 0:
            break
        ids, rp, col_ids, wts = h.partition(0, p)
        assert wts is None and ids.size == 100 and rp[-1] == 100 and np.array_equal(col_ids, (ids + 1) % n)
        seen.extend(ids.tolist())
    assert seen == list(range(n)) and h.next_partition(0) == 0      # the cycle restarts
    h.export_partition_files(0, str(tmp_path / "parts"))
    assert (tmp_path / "parts" / "partition_9").read_text().split()[:2] == ["900", "901"]
    # the reference's wrapper names
    wpr = GraphGpuWrapper()
    wpr.set_up_types(["u2i"], ["user", "item"])
    ef = tmp_path / "e.txt"
    ef.write_text("".join(f"{i}\t{(i + 1) % 50}\n" for i in range(50)))
    wpr.load_edge_file("u2i", str(ef))
    wpr.table.build(0)
    assert wpr.load_node_file("user", str(nf)) == 3 and wpr.get_all_feature_ids("user").tolist() == [7001, 7002, 8001, 8002]
    assert wpr.get_feature_of_nodes("user", np.array([1], np.uint64), [102]).tolist() == [[8001]]
    assert wpr.make_partitions("u2i", 32 * 10, 1) == 5 and wpr.get_partition_num("u2i") == 5
    assert wpr.get_partition("u2i", 1).tolist() == list(range(10, 20))
    p0 = wpr.load_next_partition("u2i", device="cpu")
    assert p0 == 0 and wpr.resident_partition == 0
    dg = wpr.device_graph
    nb, cnt = dg.sample_neighbors(torch.tensor([3, 9], dtype=torch.int64), 2, 0)
    assert cnt.tolist() == [1, 0] and int(nb[0, 0]) == 4       # 9 -> 10 leaves the resident partition: dropped


def test_graph_wrapper_service_names_and_complementary_graph(tmp_path):
    """The rest of GraphGpuWrapper's surface: feature conf, load_node_and_edge over directories (part_num,
    reverse -> inverse type), query_node_list paging, search levels, and the complementary graph of hot
    neighbours next to a resident partition."""
    import numpy as np
    from paddlebox_b200.graph import GraphGpuWrapper
    w = GraphGpuWrapper()
    w.set_up_types(["u2i", "i2u"], ["u", "i"])
    w.set_device([0])
    w.init_search_level(2)
    assert w.add_table_feat_conf("u", "age", "int64", 1) == 0 and w.add_table_feat_conf("u", "city", "int64", 1) == 1
    assert w.add_table_feat_conf("u", "age", "float32", 2) == 0 and w.feature_slot("u", "city") == 1
    assert w.add_table_feat_conf("nobody", "x", "int64", 1) == -1
    w.init_service()
    assert GraphGpuWrapper.inverse_etype("u2i") == "i2u" and GraphGpuWrapper.inverse_etype("plain") == "plain"
    # data: users 1..40 -> items 1000+; item 1000 is by far the most referenced
    ed = tmp_path / "edges" / "u2i"
    ed.mkdir(parents=True)
    rng = np.random.default_rng(0)
    n_edges = 0
    for part in range(3):
        with open(ed / f"part-{part}", "w") as f:
            for u in range(1 + part * 14, 1 + (part + 1) * 14):
                for it in {1000} | set(int(x) for x in 1001 + rng.integers(0, 30, 3)):
                    f.write(f"{u}\t{it}\n")
                    n_edges += 1
    nd = tmp_path / "nodes"
    nd.mkdir()
    with open(nd / "n0", "w") as f:
        for u in range(1, 43):
            f.write(f"u\t{u}\t0:{u % 7}\t1:{u % 3}\n")
        for it in range(1000, 1031):
            f.write(f"i\t{it}\t0:{it % 5}\n")
    got = w.load_node_and_edge("u2i", "u,i", str(tmp_path / "edges"), str(nd), part_num=0, reverse=True)
    assert got["u2i"] == n_edges and got["u"] == 42 and got["i"] == 31
    w.upload_batch(device="cpu")
    users = w.get_all_id("u2i")
    assert 1 in users.tolist() and 42 in users.tolist() or users.size >= 42 - 1
    # paging through a type's id list
    page = w.query_node_list(0, "u2i", 5, 10)
    assert np.array_equal(page, users[5:15]) and w.query_node_list(0, "u", 0, 3).tolist() == [1, 2, 3]
    # the reversed type really is the inverse relation
    nb, cnt = w.neighbor_sample("i2u", np.array([1000], np.uint64), 64)
    assert cnt[0] == 42 and set(nb[0, :42].tolist()) == set(range(1, 43))
    # part_num limits the files read
    w2 = GraphGpuWrapper()
    w2.set_up_types(["u2i"], [])
    assert 0 < w2.load_node_and_edge("u2i", "", str(tmp_path / "edges"), str(nd), part_num=1)["u2i"] < n_edges
    # complementary graph on a homogeneous relation: every node points at the hubs 1..3 and at a few others
    ef = tmp_path / "n2n.txt"
    with open(ef, "w") as f:
        for a in range(1, 61):
            for b in {1, 2, 3} | set(int(x) for x in 1 + rng.integers(0, 60, 2)):
                if a != b:
                    f.write(f"{a}\t{b}\n")
    w3 = GraphGpuWrapper()
    w3.set_up_types(["n2n"], [])
    w3.init_search_level(2)
    w3.load_edge_file("n2n", str(ef), False)
    w3.upload_batch(device="cpu")
    nparts = w3.make_partitions("n2n", 512, 1)
    assert nparts >= 3
    ids, rp, col = w3.table.csr(w3.edge_types["n2n"])
    deg = dict(zip(ids.tolist(), np.diff(rp).tolist()))
    # page in a partition that does not hold the hubs: they are what the complementary graph must bring along
    p = w3.load_next_partition("n2n", device="cpu")
    while {1, 2, 3} & set(w3.get_partition("n2n", p).tolist()):
        p = w3.load_next_partition("n2n", device="cpu")
        assert p >= 0
    resident = set(w3.get_partition("n2n", p).tolist())
    budget = 8 * (deg[1] + deg[2] + deg[3] + 2)
    chosen = w3.make_complementary_graph("n2n", budget, device="cpu")
    assert set(chosen[:3].tolist()) == {1, 2, 3}                               # most referenced first
    assert not (set(chosen.tolist()) & resident)                               # never what is resident already
    assert sum(deg[int(c)] for c in chosen) * 8 <= budget                      # fits the byte budget
    cg = w3.complementary_graph
    assert cg is not None
    # no partition resident: the hubs still lead
    w3.resident_partition = -1
    top = w3.make_complementary_graph("n2n", 1 << 20, device="cpu")
    assert set(top[:3].tolist()) == {1, 2, 3} and 40 < top.size <= ids.size     # (only referenced nodes are candidates)
    assert w3.make_complementary_graph("n2n", 0, device="cpu").size == 0
