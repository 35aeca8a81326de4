"""Framework-level pieces that run anywhere: flags, dense checkpoint format, metric variants,
CTR op family oracles, side tables, file manager, BoxPSOptimizer, data generator."""
import os
import struct

import numpy as np
import pytest
import torch

from paddlebox_b200 import checkpoint as ckpt
from paddlebox_b200.flags import FLAGS, get_flags, set_flags
from paddlebox_b200.metrics import create_metric
from paddlebox_b200.ops import ctr_ops as O
from paddlebox_b200.ops import functional as F
from paddlebox_b200.ops import reference as R
from paddlebox_b200.ops.seqpool import fused_seqpool_cvm, split_slots
from paddlebox_b200.side_tables import BoxFileMgr, GpuReplicaCache, InputTable


def test_flags_env_and_setters(monkeypatch):
    assert FLAGS.enable_pullpush_dedup_keys is True
    set_flags({"FLAGS_padbox_record_pool_max_size": "123", "check_nan_inf": 1})
    assert get_flags(["padbox_record_pool_max_size", "FLAGS_check_nan_inf"]) == {
        "padbox_record_pool_max_size": 123, "FLAGS_check_nan_inf": True}
    FLAGS.check_nan_inf = False
    with pytest.raises(AttributeError):
        FLAGS.no_such_flag = 1
    assert "gpugraph_hbm_table_load_factor" in FLAGS.describe()


def test_lod_tensor_format_roundtrip_and_layout(tmp_path):
    t = torch.arange(12, dtype=torch.float32).reshape(3, 4)
    raw = ckpt.serialize_lod_tensor(t)
    # u32 version | u64 lod levels | u32 tensor version | i32 desc size | desc | data
    assert struct.unpack_from("<I", raw, 0)[0] == 0 and struct.unpack_from("<Q", raw, 4)[0] == 0
    dsize = struct.unpack_from("<i", raw, 16)[0]
    desc = raw[20:20 + dsize]
    assert desc == b"\x08\x05\x10\x03\x10\x04"          # data_type FP32(5), dims 3, 4
    assert raw[20 + dsize:] == t.numpy().tobytes()
    back, lod = ckpt.deserialize_lod_tensor(raw)
    assert torch.equal(back, t) and lod == []
    b16 = torch.randn(5, 2).to(torch.bfloat16)
    assert torch.equal(ckpt.deserialize_lod_tensor(ckpt.serialize_lod_tensor(b16))[0], b16)
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.parallel.dense import DenseParamStore
    m = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(8,))
    store = DenseParamStore(m)
    store.m.normal_()
    names = ckpt.save_persistables(m, str(tmp_path / "dense"), store)
    assert "fc_0.w_0" in names and "fc_0.b_0" in names and "fc_0.w_0_moment1_0" in names
    m2 = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(8,))
    store2 = DenseParamStore(m2)
    ckpt.load_persistables(m2, str(tmp_path / "dense"), store2)
    for (n1, p1), (n2, p2) in zip(m.named_dense_params(), m2.named_dense_params()):
        assert n1 == n2 and torch.equal(p1, p2)
    for p, off in zip(store.params, store.offsets):      # (padding words between tensors are not saved)
        assert torch.equal(store.m[off:off + p.numel()], store2.m[off:off + p.numel()])
    # asymmetric multi-layer tower: parameter order of the store (w0, w1, .., b0, b1, ..) differs
    # from the checkpoint's name order (w0, b0, w1, b1, ..); moments must follow their parameter, and FC
    # weights are stored in the reference's [in, out] layout
    m3 = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(16, 8))
    s3 = DenseParamStore(m3)
    s3.m.normal_(); s3.v.uniform_()
    ckpt.save_persistables(m3, str(tmp_path / "d3"), s3)
    w0 = dict(m3.named_dense_params())["fc_0.w_0"]
    on_disk, _ = ckpt.deserialize_lod_tensor(open(tmp_path / "d3" / "fc_0.w_0", "rb").read())
    assert tuple(on_disk.shape) == (w0.shape[1], w0.shape[0]) and torch.equal(on_disk, w0.detach().t())
    m4 = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(16, 8))
    s4 = DenseParamStore(m4)
    ckpt.load_persistables(m4, str(tmp_path / "d3"), s4)
    for (n1, p1), (n2, p2) in zip(m3.named_dense_params(), m4.named_dense_params()):
        assert n1 == n2 and torch.equal(p1, p2)
    off3 = {id(p): o for p, o in zip(s3.params, s3.offsets)}
    off4 = {id(p): o for p, o in zip(s4.params, s4.offsets)}
    for (n1, p1), (n2, p2) in zip(m3.named_dense_params(), m4.named_dense_params()):
        a, b, n = off3[id(p1)], off4[id(p2)], p1.numel()
        assert torch.equal(s3.m[a:a + n], s4.m[b:b + n]) and torch.equal(s3.v[a:a + n], s4.v[b:b + n]), n1
    # a tensor of the right size but the wrong shape is refused, not reshaped
    m5 = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(8, 16))
    import pytest as _pt
    with _pt.raises(ValueError):
        ckpt.load_persistables(m5, str(tmp_path / "d3"), strict=False)


def test_metric_variants():
    rs = np.random.RandomState(0)
    n = 2000
    pred = torch.from_numpy(rs.rand(n).astype(np.float32))
    label = (torch.from_numpy(rs.rand(n).astype(np.float32)) < pred).float()
    mask = torch.from_numpy((rs.rand(n) < 0.5).astype(np.int64))
    cm = torch.from_numpy(rs.choice([222, 223], n))
    rk = torch.from_numpy(rs.choice([1, 2, 3], n))
    cmatch_rank = (cm << 32) | rk
    t = {"pred": pred, "label": label, "mask": mask, "cmatch_rank": cmatch_rank, "m2": (mask * 2),
         "pred_b": pred * 0.5, "uid": torch.from_numpy(rs.randint(0, 50, n))}
    base = create_metric("AucCalculator", name="a", bucket_size=10000)
    base.add_data(t)
    full = base.get_metric_msg()
    assert full[7] == n and 0.6 < full[0] < 0.9
    mm = create_metric("MaskAucCalculator", name="m", mask_varname="mask", bucket_size=10000)
    mm.add_data(t)
    assert mm.get_metric_msg()[7] == int(mask.sum())
    mu = create_metric("MultiMaskAucCalculator", name="mu", mask_varname="mask m2", mask_varvalue_list="1 2",
                       bucket_size=10000)
    mu.add_data(t)
    assert mu.get_metric_msg()[7] == int(mask.sum())
    cr = create_metric("CmatchRankAucCalculator", name="c", cmatch_rank_group="222_1 223_2",
                       cmatch_rank_varname="cmatch_rank", bucket_size=10000)
    cr.add_data(t)
    exp = int((((cm == 222) & (rk == 1)) | ((cm == 223) & (rk == 2))).sum())
    assert cr.get_metric_msg()[7] == exp
    ci = create_metric("CmatchRankAucCalculator", name="ci", cmatch_rank_group="222", ignore_rank=True,
                       cmatch_rank_varname="cmatch_rank", bucket_size=10000)
    ci.add_data(t)
    assert ci.get_metric_msg()[7] == int((cm == 222).sum())
    crm = create_metric("CmatchRankMaskAucCalculator", name="crm", cmatch_rank_group="222_1", mask_varname="mask",
                        cmatch_rank_varname="cmatch_rank", bucket_size=10000)
    crm.add_data(t)
    assert crm.get_metric_msg()[7] == int(((cm == 222) & (rk == 1) & (mask != 0)).sum())
    mt = create_metric("MultiTaskAucCalculator", name="mt", pred_varname="pred pred_b",
                       cmatch_rank_group="222_1 223_1", cmatch_rank_varname="cmatch_rank", bucket_size=10000)
    mt.add_data(t)
    assert mt.get_metric_msg()[7] == int(((rk == 1)).sum())
    fm = create_metric("FloatMaskAucCalculator", name="f", mask_varname="mask", bucket_size=10000)
    fm.add_data({**t, "label": pred})     # float labels
    assert abs(fm.get_metric_msg()[7] - int(mask.sum())) < 1e-6
    cont = create_metric("ContinueMaskCalculator", name="co", mask_varname="mask", bucket_size=10000)
    cont.add_data({**t, "label": pred})
    mae, rmse, pm, am, size = cont.get_metric_msg()
    assert mae < 1e-6 and size == int(mask.sum())
    wu = create_metric("WuAucCalculator", name="w", uid_varname="uid")
    wu.add_data(t)
    uauc, wuauc, users, size = wu.get_metric_msg()
    assert users > 10 and 0.5 < wuauc < 1.0
    ni = create_metric("NanInfCalculator", name="n")
    ni.add_data({"pred": torch.tensor([0.1, float("nan"), float("inf")]), "label": torch.zeros(3)})
    assert ni.get_nan_inf_msg() == [1, 1, 3]
    with pytest.raises(ValueError):
        create_metric("NoSuchCalculator", name="x")


def test_rank_attention_matches_numpy_oracle():
    rs = np.random.RandomState(1)
    n, C, out, max_rank = 6, 4, 5, 3
    x = rs.rand(n, C).astype(np.float32)
    param = rs.rand(max_rank * max_rank * C, out).astype(np.float32)
    sid = torch.tensor([7, 7, 7, 9, 9, 11])
    rk = torch.tensor([1, 2, 3, 2, 1, 5])          # rank 5 > max_rank: ignored
    ro = O.build_rank_offset(sid, rk, max_rank)
    assert ro[0].tolist() == [1, 1, 0, 2, 1, 3, 2] and ro[5, 0] == -1
    got = O.rank_attention(torch.from_numpy(x), ro, torch.from_numpy(param), max_rank)
    # oracle restated from the reference's numpy test (test_rank_attention_op.py:25-108)
    exp = np.zeros((n, out), np.float32)
    for i in range(n):
        lower = ro[i, 0].item() - 1
        for k in range(max_rank):
            faster = ro[i, 2 * k + 1].item() - 1
            if lower < 0 or faster < 0:
                continue
            idx = ro[i, 2 * k + 2].item()
            blk = param[(lower * max_rank + faster) * C:(lower * max_rank + faster + 1) * C]
            exp[i] += x[idx] @ blk
    assert np.allclose(got.numpy(), exp, rtol=1e-5, atol=1e-6)
    xt = torch.from_numpy(x).requires_grad_(True)
    O.rank_attention(xt, ro, torch.from_numpy(param), max_rank).sum().backward()
    assert xt.grad is not None and xt.grad.abs().sum() > 0


def test_batch_fc_scaled_fc_cross_norm_and_slices():
    torch.manual_seed(0)
    x = torch.rand(4, 6, 5)
    w, b = torch.rand(4, 5, 3), torch.rand(4, 3)
    exp = torch.stack([x[s] @ w[s] + b[s] for s in range(4)])
    assert torch.allclose(O.batch_fc(x, w, b), exp, atol=1e-6)
    assert O.BatchFC(4, 5, 3)(x).shape == torch.Size([4, 6, 3])
    xi, wi, bi = torch.randn(8, 16, requires_grad=True), torch.randn(16, 4, requires_grad=True), torch.randn(4)
    y = O.scaled_fc(xi, wi, bi, input_scale_factor=2.0, bias_scale_factor=0.5)
    assert torch.allclose(y, xi @ wi + 0.25 * bi, rtol=3e-2, atol=6e-2)        # scales cancel but for bias_scale / input_scale
    y.sum().backward()
    assert torch.allclose(xi.grad, torch.ones(8, 4) @ wi.t(), rtol=3e-2, atol=3e-2)
    assert torch.allclose(wi.grad, xi.detach().t() @ torch.ones(8, 4), rtol=3e-2, atol=6e-2)
    assert torch.isnan(O.scaled_fc(torch.full((1, 2), 3e38), torch.full((2, 1), 3e38), torch.zeros(1))).all()
    # int8: integer product of the quantised operands, rescaled by the INPUT interval only (reference quirk)
    xq, wq = torch.tensor([[0.5, -0.7, 3.0]]), torch.tensor([[1.0], [0.25], [-2.5]])
    q = lambda v: torch.trunc(v.clamp(-2, 2) * 60 + 0.5)
    assert q(xq).tolist() == [[30.0, -41.0, 120.0]] and q(wq).reshape(-1).tolist() == [60.0, 15.0, -119.0]   # -119.5 truncates toward zero
    want = (q(xq) @ q(wq)) / 60.0 + 1.0
    got8 = O.scaled_int8fc(xq, wq, torch.ones(1))
    assert torch.allclose(got8, want)
    real = O.scaled_int8fc(xq, wq, torch.ones(1), dequant_weights=True)
    assert torch.allclose(real, xq.clamp(-2, 2) @ wq.clamp(-2, 2) + 1.0, atol=0.05)
    xg = xq.clone().requires_grad_(True)
    O.scaled_int8fc(xg, wq, torch.ones(1)).sum().backward()
    assert torch.equal(xg.grad, wq.t())
    cn = O.CrossNormHadamard(fields_num=2, embed_dim=3)
    xin = torch.rand(5, 12)
    out = cn(xin)
    assert out.shape == (5, 20)
    # summaries start at size 1e4, sum 0, square_sum 1e4: mean 0, scale 1 -> the raw block [a, b, a*b, <a,b>]
    a0, b0 = xin[:, 0:3], xin[:, 3:6]
    assert torch.allclose(out[:, :10], torch.cat([a0, b0, a0 * b0, (a0 * b0).sum(1, keepdim=True)], 1), atol=1e-6)
    a, b2 = torch.arange(12.).reshape(3, 4), torch.arange(12., 24.).reshape(3, 4)
    assert torch.equal(O.partial_concat([a, b2], 1, 2), torch.cat([a[:, 1:3], b2[:, 1:3]], 1))
    assert torch.equal(O.partial_sum([a, b2], 1, 2), a[:, 1:3] + b2[:, 1:3])
    assert torch.equal(O.fused_concat([a, b2], 2, -1), torch.cat([a[:, 2:], b2[:, 2:]], 1))
    sh, idx = O.shuffle_batch(a, seed=3)
    assert torch.equal(sh, a[idx]) and sorted(idx.tolist()) == [0, 1, 2]
    kept, imap, w8 = O.filter_by_instag(a, [0, 1, 3, 4], torch.tensor([5, 6, 7, 8]), [7])
    assert torch.equal(kept, a[1:2]) and imap.tolist() == [[0, 1, 1]]
    none, nmap, lw = O.filter_by_instag(a, [0, 1, 3, 4], torch.tensor([5, 6, 7, 8]), [99])
    assert none.shape == (1, 4) and lw.item() == 0 and nmap.tolist() == [[0, 1, 1]]
    # the reference's first test case (test_filter_by_instag_op.py:30-72): instance i owns i+1 rows
    x1 = torch.arange(36.)[:, None].repeat(1, 4).requires_grad_(True)
    lod1 = [0, 1, 3, 6, 10, 15, 21, 28, 36]
    out, mm, lw = O.filter_by_instag(x1, list(range(9)), torch.tensor([1, 2, 1, 2, 1, 2, 1, 2]), [2], ins_lod=lod1)
    assert mm.tolist() == [[0, 1, 2], [2, 6, 4], [6, 15, 6], [12, 28, 8]] and out.lod == [0, 2, 6, 12, 20]
    assert out[:, 0].tolist() == [1, 2, 6, 7, 8, 9] + list(range(15, 21)) + list(range(28, 36)) and lw.tolist() == [[1.]] * 4
    out.sum().backward()
    assert x1.grad[:, 0].tolist() == [0, 1, 1, 0, 0, 0, 1, 1, 1, 1] + [0] * 5 + [1] * 6 + [0] * 7 + [1] * 8


def test_seqpool_autograd_wrapper_and_variants_cpu():
    gen = torch.Generator().manual_seed(0)
    B, S, H = 5, 3, 7
    lens = torch.randint(0, 4, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    x = (torch.rand(int(lod[-1]), H, generator=gen) * 3).requires_grad_(True)
    cvm = torch.rand(B, 2, generator=gen)
    out = fused_seqpool_cvm(x, cvm, lod, B, S)
    assert out.shape == (B, S * H) and len(split_slots(out, S)) == S
    out.sum().backward()
    # gradient: first two columns carry the cvm input, the rest ones
    g = x.grad
    assert torch.all(g[:, 2:] == 1)
    b_of_k = torch.repeat_interleave(torch.arange(S * B) % B, lens)
    assert torch.allclose(g[:, :2], cvm[b_of_k])
    # PCOC: 7 statistics columns (show, clk, show2, clk2, 2 pclk, 1 unused), checked bag by bag
    H2, used, mx = 7 + 3, 6, 7
    xp = (torch.rand(int(lod[-1]), H2, generator=gen) * 3).requires_grad_(True)
    cvm4, qv = torch.rand(B, 4, generator=gen), torch.rand(B, 2, generator=gen)
    pc = O.fused_seqpool_cvm_with_pcoc(xp, lod, B, S, cvm=cvm4, q_values=qv, cvm_offset=used, max_cvm_offset=mx)
    Eo = 2 + 2 * 2 + 3
    assert pc.shape == (B, S * Eo)
    ll = lod.tolist()
    for s_ in range(S):
        for b_ in range(B):
            p_ = xp.detach()[ll[s_ * B + b_]:ll[s_ * B + b_ + 1]].sum(0)
            lg = torch.log(p_ + 1)
            want = torch.cat([lg[0:1], lg[1:2] - lg[0:1], lg[4:6] - lg[2], lg[4:6] - lg[3], p_[7:]])
            assert torch.allclose(pc[b_, s_ * Eo:(s_ + 1) * Eo], want, atol=1e-5)
    pc.sum().backward()
    gp = xp.grad
    assert torch.allclose(gp[:, :4], cvm4[b_of_k]) and torch.allclose(gp[:, 4:6], qv[b_of_k])
    assert torch.all(gp[:, 6] == 0) and torch.all(gp[:, 7:] == 1)
    nocvm = O.fused_seqpool_cvm_with_pcoc(xp.detach(), lod, B, S, use_cvm=False, cvm_offset=used, max_cvm_offset=mx)
    assert nocvm.shape == (B, S * 3)


def test_side_tables_and_file_mgr(tmp_path):
    c = GpuReplicaCache(4)
    for i in range(5):
        assert c.add_items([i, i + 1, i + 2, i + 3]) == i
    c.to_hbm(torch.device("cpu"))
    got = c.pull_cache_value(torch.tensor([4, 0]))
    assert got.tolist() == [[4, 5, 6, 7], [0, 1, 2, 3]]
    t = InputTable(2)
    t.add_index_data("ad_1", [1.0, 2.0])
    idxf = tmp_path / "idx.txt"
    idxf.write_text("ad_2\t3 4\nbad\t1\n")
    assert t.load_index_file(str(idxf)) == 1
    offs = torch.tensor([t.get_index_offset("ad_2"), t.get_index_offset("nope"), t.get_index_offset("ad_1")])
    assert t.lookup_input(offs).tolist() == [[3, 4], [0, 0], [1, 2]] and t.miss == 1
    fm = BoxFileMgr()
    d = str(tmp_path / "fs")
    assert fm.makedir(d) and fm.exists(d)
    fm.touch(os.path.join(d, "a"))
    open(os.path.join(d, "b"), "w").write("hello")
    assert fm.count(d) == 2 and fm.file_size(os.path.join(d, "b")) == 5
    fm.rename(os.path.join(d, "b"), os.path.join(d, "c"))
    fm.truncate(os.path.join(d, "c"), 2)
    fm.upload(os.path.join(d, "c"), os.path.join(d, "e"))
    assert sorted(os.path.basename(p) for p in fm.list_dir(d)) == ["a", "c", "e"]
    assert dict((os.path.basename(p), s) for p, s in fm.dus(d))["e"] == 2
    fm.remove(d)
    assert not fm.exists(d)


def test_boxps_optimizer_and_data_generator(tmp_path):
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.optimizer import Adam, BoxPSOptimizer, LocalSGD
    m = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(8,))
    cfg = BoxPSOptimizer(Adam(learning_rate=0.01), sync_steps=4, sync_dense_mode=1).minimize(m)
    assert cfg.dense_lr == 0.01 and cfg.sync_dense_mode == "kstep" and cfg.sync_weight_step == 4
    assert m._pipeline_opt is cfg and "tower.weights.0" in m._param_need_sync
    assert LocalSGD(k_steps=8).transpile(main_program=m) == "kstep" and cfg.sync_weight_step == 8
    from paddlebox_b200.data_generator import MultiSlotDataGenerator

    class G(MultiSlotDataGenerator):
        def generate_sample(self, line):
            def it():
                a, b = line.split()
                yield [("label", [float(a)]), ("slot", [int(b), int(b) + 1])]
            return it
    out = tmp_path / "gen.txt"
    with open(out, "w") as f:
        G().run_from_iterable(["1 5", "0 9"], f)
    assert out.read_text() == "1 1.0 2 5 6\n1 0.0 2 9 10\n"


def test_data_norm_module_updates_summaries():
    dn = F.DataNorm(3, summary_decay_rate=0.5)
    x = torch.tensor([[1., 2., 3.], [3., 2., 1.]], requires_grad=True)
    y = dn(x)
    y.sum().backward()
    assert torch.allclose(dn.batch_size, torch.full((3,), 1e4 * 0.5 + 1))
    assert torch.allclose(dn.batch_sum, torch.tensor([2., 2., 2.]))
    assert torch.allclose(x.grad, torch.ones(2, 3) * torch.sqrt(torch.tensor(1e4 / 1e4)))


def test_input_table_dataset_loads_index_files(tmp_path):
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory
    from tests.util_data import slot_vars, write_files
    BoxWrapper.reset_instance()
    box = BoxWrapper.set_instance(embedx_dim=8, device=torch.device("cpu"), rank=0, world_size=1)
    box.initialize_gpu_and_load_model({})
    try:
        idx = tmp_path / "index-0"
        idx.write_text("query_a\t0.5 1.5 2.5\nquery_b\t1 2 3\nbad line\n")
        ds = DatasetFactory().create_dataset("InputTableDataset")
        ds.set_use_var(slot_vars(3, 2))
        ds.set_batch_size(8)
        ds.set_thread(1)
        ds.set_input_table_dim(3)
        ds.set_index_filelist([str(idx)])
        ds.set_filelist(write_files(str(tmp_path), n_files=1, lines_per_file=16, n_sparse=3, n_dense=2, key_space=20))
        ds.load_into_memory()
        assert ds.input_table.size() == 2 and ds.get_memory_data_size() == 16
        off = ds.index_offsets(["query_b", "nope", "query_a"])
        assert off.tolist() == [2, 0, 1] and ds.input_table.miss == 1
        rows = ds.input_table.lookup_input(off)
        assert torch.allclose(rows, torch.tensor([[1., 2., 3.], [0., 0., 0.], [0.5, 1.5, 2.5]]))
    finally:
        BoxWrapper.reset_instance()


def test_parser_plugin_resolves_input_table_keys_at_parse_time(tmp_path):
    """GetOffsetFunc: a plugin that exports pbx_set_offset_func turns the string key of every line into
    the InputTable row offset while parsing; lookup_input then gathers the rows."""
    import subprocess
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory, dense_slot, sparse_slot
    src = tmp_path / "plug.cc"
    src.write_text(r'''
#include <cstdint>
#include <cstdlib>
#include <cstring>
static uint64_t (*g_off)(void*, const char*, size_t) = nullptr;
static void* g_user = nullptr;
extern "C" void pbx_set_offset_func(uint64_t (*fn)(void*, const char*, size_t), void* user) { g_off = fn; g_user = user; }
extern "C" int pbx_parse_line(const char* line, size_t len, void* ctx,
    void (*emit_u64)(void*, int, const uint64_t*, size_t), void (*emit_f32)(void*, int, const float*, size_t),
    void (*emit_meta)(void*, const char*, uint64_t, uint32_t, uint32_t)) {
  // format: "<label> <string key> <feasign>"
  char* end = nullptr;
  float label = strtof(line, &end);
  const char* k = end + 1;
  const char* ke = (const char*)memchr(k, ' ', line + len - k);
  if (!ke || !g_off) return 1;
  uint64_t off = g_off(g_user, k, (size_t)(ke - k));
  uint64_t fea = strtoull(ke + 1, &end, 10);
  emit_f32(ctx, 0, &label, 1);
  emit_u64(ctx, 1, &off, 1);          // the offset; 0 (miss) is dropped like any zero feasign
  emit_u64(ctx, 2, &fea, 1);
  emit_meta(ctx, "", 0, 0, 0);
  return 0;
}''')
    so = tmp_path / "plug.so"
    subprocess.check_call(["g++", "-shared", "-fPIC", "-O1", "-o", str(so), str(src)])
    BoxWrapper.reset_instance()
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    box.initialize_gpu_and_load_model({}, slot_omit_in_feedpass=["input_off"])
    try:
        (tmp_path / "index-0").write_text("qa\t1 2\nqb\t3 4\n")
        (tmp_path / "d.txt").write_text("1 qb 11\n0 zz 12\n1 qa 13\n")
        ds = DatasetFactory().create_dataset("InputTableDataset")
        ds.set_use_var([dense_slot("label"), sparse_slot("input_off"), sparse_slot("a")])
        ds._native.set("parser_plugin", str(so))
        ds.set_batch_size(3)
        ds.set_thread(1)
        ds.set_input_table_dim(2)
        ds.set_index_filelist([str(tmp_path / "index-0")])
        ds.set_filelist([str(tmp_path / "d.txt")])
        ds.load_into_memory()
        assert ds.get_memory_data_size() == 3 and ds.input_table.miss == 1
        ds._native.prepare_train(1, False, 0, 0)
        # the offset slot is named in slot_omit_in_feedpass: its values are not parameter-server keys
        assert sorted(ds.unique_keys().tolist()) == [0, 11, 12, 13]
        assert sorted(box._pass_keys.tolist()) == [0, 11, 12, 13]
        rows = ds.input_table.lookup_input(torch.tensor([2, 0, 1]))
        assert torch.allclose(rows, torch.tensor([[3., 4.], [0., 0.], [1., 2.]]))
    finally:
        BoxWrapper.reset_instance()


def test_queue_dataset_streams_in_file_order(tmp_path):
    import torch
    from paddlebox_b200.dataset import DatasetFactory
    from tests.util_data import slot_vars, write_files
    ds = DatasetFactory().create_dataset("QueueDataset")
    ds.set_use_var(slot_vars(2, 1))
    ds.set_batch_size(5)
    ds.set_thread(1)
    ds.set_merge_thread_num(1)
    files = write_files(str(tmp_path), n_files=1, lines_per_file=20, n_sparse=2, n_dense=1, key_space=1000)
    ds.set_filelist(files)
    ds.load_into_memory()
    with pytest.raises(NotImplementedError):
        ds.local_shuffle()
    ds.prepare_train(1)
    labels = torch.cat([b.dense[:, 0] for b in ds.batches(device=torch.device("cpu"))])
    # file order is preserved: the labels come back in the order they were written
    want = [float(line.split()[1]) for line in open(files[0])]
    assert labels.tolist() == want
    assert ds.unique_keys().size == 0            # a streaming dataset does not collect pass keys


def test_vectorised_cpu_seqpool_matches_the_loop_oracle():
    import torch
    from paddlebox_b200.ops import reference as R
    gen = torch.Generator().manual_seed(0)
    B, S, H = 7, 5, 9
    lens = torch.randint(0, 4, (S * B,), generator=gen)
    lod = torch.zeros(S * B + 1, dtype=torch.int32)
    lod[1:] = torch.cumsum(lens, 0)
    x = torch.rand(int(lod[-1]), H, generator=gen) * 3
    cvm = torch.rand(B, 4, generator=gen)
    for kw in (dict(use_cvm=True, cvm_offset=2), dict(use_cvm=True, cvm_offset=2, clk_filter=True),
               dict(use_cvm=False, cvm_offset=2), dict(use_cvm=True, cvm_offset=3, variant=1),
               dict(use_cvm=True, cvm_offset=3, variant=1, show_filter=True), dict(use_cvm=True, cvm_offset=4, variant=2),
               dict(use_cvm=False, cvm_offset=3, variant=1), dict(use_cvm=True, cvm_offset=2, pad_value=0.5)):
        assert R.seqpool_plain(kw)
        a = R.seqpool_cvm_fwd(x, lod, B, S, **kw)
        b = R.seqpool_cvm_fwd_vec(x, lod, B, S, **kw)
        assert a.shape == b.shape and torch.allclose(a, b, atol=1e-6), kw
        dout = torch.rand(B, a.shape[1], generator=gen)
        bkw = {k: v for k, v in kw.items() if k != "pad_value"}
        da = R.seqpool_cvm_bwd(dout, cvm, lod, B, S, H, **bkw)
        db = R.seqpool_cvm_bwd_vec(dout, cvm, lod, B, S, H, **bkw)
        assert torch.equal(da, db), kw
    assert not R.seqpool_plain(dict(need_filter=True)) and not R.seqpool_plain(dict(embedx_concate_size=2))


def test_float_label_metrics_accumulate_without_host_sync():
    """The device-side accumulation used for float-label / continuous metrics (pure PyTorch ops, so it is checked
    here on CPU tensors) gives the numbers of the host calculator (csrc/host/metrics.cc add_data)."""
    from paddlebox_b200.host import HostAuc
    from paddlebox_b200.metrics import accumulate_float_labels
    rs = np.random.RandomState(3)
    n, buckets = 5000, 1000
    pred = torch.from_numpy(rs.rand(n).astype(np.float32))
    pred[:3] = torch.tensor([float("nan"), 1.7, -0.2])
    label = torch.from_numpy(rs.rand(n).astype(np.float32))
    mask = torch.from_numpy((rs.rand(n) < 0.7).astype(np.float32))
    for mode in (1, 2):
        ref = HostAuc(buckets)
        ref.add(pred.numpy(), label.numpy(), mask.numpy(), mode)
        rt, rsum = ref.export_state()
        table = torch.zeros(2, buckets, dtype=torch.float64)
        sums = torch.zeros(5, dtype=torch.float64)
        accumulate_float_labels(table, sums, pred[:2000], label[:2000], mask[:2000], mode)       # two batches
        accumulate_float_labels(table, sums, pred[2000:], label[2000:], mask[2000:], mode)
        if mode == 1:
            assert np.allclose(table.numpy(), rt, atol=1e-9)
            assert np.allclose(sums.numpy(), rsum, rtol=1e-9)
        else:                                   # continuous: NaN is not clamped by either side
            ok = ~torch.isnan(pred)
            ref2 = HostAuc(buckets)
            ref2.add(pred[ok].numpy(), label[ok].numpy(), mask[ok].numpy(), 2)
            t2 = torch.zeros(2, buckets, dtype=torch.float64); s2 = torch.zeros(5, dtype=torch.float64)
            accumulate_float_labels(t2, s2, pred[ok], label[ok], mask[ok], 2)
            assert np.allclose(s2.numpy(), ref2.export_state()[1], rtol=1e-9) and float(t2.sum()) == 0.0


def test_sorted_merge_matches_atomic_merge_under_key_skew():
    """Hot-key batch: 30 % of the occurrences are one key, plus the padding key 0 — the sorted fp64 merge
    (reference PushMergeCopy) and the index_add merge agree, and the sorted one is bit-reproducible."""
    from paddlebox_b200.ops import reference as R
    g = torch.Generator().manual_seed(9)
    n, H, S = 6000, 11, 5
    keys = torch.randint(1, 400, (n,), generator=g)
    keys[torch.rand(n, generator=g) < 0.30] = 77          # the hot key
    keys[:50] = 0                                         # padding key
    uk, restore = torch.unique(keys, return_inverse=True)
    grads = torch.randn(n, H, generator=g)
    grads[:, 0] = 1.0
    grads[:, 1] = (torch.rand(n, generator=g) < 0.2).float()
    k2s = torch.randint(0, S, (n,), generator=g)
    a = R.merge_push(restore, uk.numel(), grads, k2s, 2, -64.0)
    b = R.merge_push_sorted(restore, uk.numel(), grads, k2s, 2, -64.0)
    assert torch.allclose(a[:, 1:], b[:, 1:], rtol=1e-5, atol=1e-4)
    hot = int((uk == 77).nonzero())
    assert abs(float(b[hot, 1]) - float((keys == 77).sum())) < 1e-3          # show count of the hot key
    perm = torch.randperm(n, generator=g)                                     # another occurrence order
    c = R.merge_push_sorted(restore[perm], uk.numel(), grads[perm], k2s[perm], 2, -64.0)
    assert torch.allclose(b[:, 1:], c[:, 1:], rtol=1e-6, atol=1e-5)
    again = R.merge_push_sorted(restore, uk.numel(), grads, k2s, 2, -64.0)
    assert torch.equal(b, again)                                               # bit-reproducible
    first = torch.tensor([int((restore == u).nonzero()[0]) for u in range(uk.numel())])
    assert torch.equal(b[:, 0], k2s[first].float())                            # slot of the first occurrence


def test_cpu_training_with_deterministic_merge_flag(tmp_path):
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.flags import FLAGS
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch
    from paddlebox_b200.trainer import CTRTrainer

    def run(flag):
        BoxWrapper.reset_instance()
        FLAGS.pbx_deterministic_merge = flag
        try:
            box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
            box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
            g = torch.Generator().manual_seed(1)
            keys = torch.randint(1, 30, (96,), generator=g)
            keys[::3] = 7
            lod = torch.arange(0, 97, 2, dtype=torch.int32)          # 48 bags of 2 keys: 2 slots x 24 instances
            dense = torch.rand(24, 2, generator=g); dense[:, 0] = (dense[:, 0] < 0.3).float()
            b = SlotBatch(keys, lod, 24, 2, dense)
            box.begin_feed_pass("20240101"); box.feed_pass(keys); box.end_feed_pass(); box.begin_pass()
            torch.manual_seed(0)
            tr = CTRTrainer(CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(8,)), box)
            losses = [float(tr.train_step(b).detach()) for _ in range(6)]
            box.end_pass()
            return losses
        finally:
            FLAGS.pbx_deterministic_merge = False
            BoxWrapper.reset_instance()
    a, b = run(False), run(True)
    assert np.allclose(a, b, rtol=1e-4, atol=1e-5) and run(True) == b


def test_bf16_tower_numerics_train_to_the_same_auc_as_fp32(tmp_path):
    """SURVEY 7.3 asks for a check that bf16 towers (fp32 accumulation, fp32 master weights) train like the
    reference's fp32 towers.  The GPU numerics are stated on the CPU (FLAGS.pbx_cpu_emulate_bf16): same data,
    same seeds, DeepFM, AUC per pass and final loss must agree closely — and the runs must really differ."""
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.flags import set_flags
    from paddlebox_b200.models.ctr import DeepFM
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    from tests.util_data import slot_vars, write_files
    files = write_files(str(tmp_path), n_files=2, lines_per_file=400, n_sparse=8, n_dense=4, key_space=40)

    def run(emulate):
        set_flags({"pbx_cpu_emulate_bf16": emulate})
        BoxWrapper.reset_instance()
        box = BoxWrapper.set_instance(embedx_dim=8, device=torch.device("cpu"), rank=0, world_size=1)
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 1.0, "learning_rate": 0.1, "mf_learning_rate": 0.1})
        box.init_metric("AucCalculator", "auc", "label", "pred", metric_phase=1, bucket_size=100000)
        torch.manual_seed(0)
        ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
        ds.set_use_var(slot_vars(8, 4)); ds.set_batch_size(32); ds.set_thread(1); ds.set_filelist(files)
        ds.set_merge_thread_num(1)                      # one reader, one merger: the record order is the file order
        model = DeepFM(num_slots=8, embedx_dim=8, dense_dim=4, hidden=(64, 32))
        tr = CTRTrainer(model, box, TrainerConfig(dense_lr=1e-2))
        aucs, loss = [], None
        for _ in range(8):
            ds.load_into_memory(); ds.begin_pass()
            st = tr.train_from_dataset(ds)
            ds.end_pass(False)
            aucs.append(box.get_metric_msg("auc")[0])
            loss = st.get("loss", st.get("mean_loss"))
            ds.release_memory()
        flat = tr.store.flat.clone()
        BoxWrapper.reset_instance()
        return aucs, loss, flat

    try:
        a32, l32, p32 = run(False)
        a16, l16, p16 = run(True)
    finally:
        set_flags({"pbx_cpu_emulate_bf16": False})
    assert not torch.equal(p32, p16)                      # the emulation really changed the arithmetic
    assert a32[-1] > 0.7 and a16[-1] > 0.7
    assert abs(a32[-1] - a16[-1]) < 0.01, (a32, a16)
    assert max(abs(x - y) for x, y in zip(a32, a16)) < 0.02, (a32, a16)


def test_queue_dataset_streams_without_loading_the_pass(tmp_path):
    """No load_into_memory: files are read a group at a time while the previous group's batches are consumed;
    every line arrives exactly once, in file order with one reader, and at most two groups are resident."""
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    from tests.util_data import slot_vars, write_files
    files = write_files(str(tmp_path), n_files=5, lines_per_file=37, n_sparse=3, n_dense=2, key_space=50)
    ds = DatasetFactory().create_dataset("QueueDataset")
    ds.set_use_var(slot_vars(3, 2)); ds.set_batch_size(8); ds.set_thread(1); ds.set_merge_thread_num(1)
    ds.set_filelist(files)
    assert not ds.is_loaded() and ds.file_groups() == [[f] for f in files]
    labels = torch.cat([b.dense[:, 0] for b in ds.batches(device=torch.device("cpu"))])
    want = [float(line.split()[1]) for f in files for line in open(f)]
    assert labels.tolist() == want                                   # every line once, in file order
    assert ds.stream_stats["records"] == 5 * 37 and ds.stream_stats["groups"] == 5
    assert ds.stream_stats["max_resident_records"] <= 2 * 37          # two groups at most, never the pass
    assert ds.get_memory_data_size() == 0                             # the dataset itself holds nothing
    fi = DatasetFactory().create_dataset("FileInstantDataset")
    fi.set_use_var(slot_vars(3, 2)); fi.set_batch_size(8); fi.set_thread(1); fi.set_filelist(files[:2])
    with pytest.raises(NotImplementedError):
        fi.load_into_memory()
    assert sum(b.batch_size for b in fi.batches(device=torch.device("cpu"))) == 2 * 37
    # several files per group with several readers: nothing lost, nothing doubled
    ds.set_thread(2)
    assert [len(g) for g in ds.file_groups()] == [2, 2, 1]
    n = sum(b.batch_size for b in ds.batches(device=torch.device("cpu")))
    assert n == 5 * 37
    # a bad file in a later group surfaces as an error in the consumer, not as a silent short pass
    ds.set_filelist(files[:2] + [str(tmp_path / "missing")])
    ds.set_thread(1)
    with pytest.raises(Exception):
        for _ in ds.batches(device=torch.device("cpu")):
            pass
    # and it trains through the public entry point without a feed pass: keys are created on first sight
    ds.set_filelist(files)
    BoxWrapper.reset_instance()
    try:
        box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
        tr = CTRTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=2, hidden=(8,)), box, TrainerConfig())
        box.begin_pass()
        st = tr.train_from_dataset(ds)
        box.end_pass()
        assert st["instances"] == 5 * 37 and st["steps"] >= 5 * 5
    finally:
        BoxWrapper.reset_instance()


def test_fetch_handler_is_called_periodically_during_training(tmp_path):
    import numpy as np
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, Executor, FetchHandler, TrainerConfig
    from tests.util_data import slot_vars, write_files
    files = write_files(str(tmp_path), n_files=1, lines_per_file=600, n_sparse=3, n_dense=2, key_space=50)
    BoxWrapper.reset_instance()
    try:
        box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
        ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
        ds.set_use_var(slot_vars(3, 2)); ds.set_batch_size(4); ds.set_thread(1); ds.set_filelist(files)
        ds.load_into_memory(); ds.begin_pass()
        tr = CTRTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=2, hidden=(8,)), box, TrainerConfig())
        seen = []

        class H(FetchHandler):
            def handler(self, res):
                seen.append(res)
                if len(seen) == 2:
                    raise RuntimeError("a failing handler must not stop the pass")

        st = Executor().train_from_dataset(tr, ds, fetch_handler=H({"p": "pred", "steps": lambda: tr.steps,
                                                                   "nothing": "no_such_output"}, period_secs=0.02))
        ds.end_pass(False)
        assert st["instances"] == 600 and len(seen) >= 2
        last = seen[-1]
        assert isinstance(last["p"], np.ndarray) and last["p"].size == 4 and last["nothing"] is None
        assert int(last["steps"]) > 0
        FetchHandler({"a": 1}).handler({"a": np.array([3.0])})
        FetchHandler.help()
    finally:
        BoxWrapper.reset_instance()


def test_run_cmd_and_stage_timers(tmp_path):
    """``core.run_cmd`` of the reference's test_boxps.py (:96-106): output of a shell command, with retries
    of a failing one until the time-out; and the host stage timers behind ``print_sync_timer``."""
    import time
    from paddlebox_b200.utils.shell import run_cmd
    from paddlebox_b200.utils.timer import CudaStageTimers, StageTimers
    assert run_cmd("echo hello").strip() == "hello"
    assert run_cmd("ls /no/such/dir/at/all 2>/dev/null") == ""
    flag = tmp_path / "flag"
    # fails twice (no file), then succeeds: retried every 20 ms within the 2 s time-out
    cmd = f"if [ -f {flag}.2 ]; then echo done; elif [ -f {flag}.1 ]; then touch {flag}.2; exit 1; else touch {flag}.1; exit 1; fi"
    t0 = time.time()
    assert run_cmd(cmd, time_out=2000, sleep_inter=20).strip() == "done" and time.time() - t0 < 2.0
    t0 = time.time()
    assert run_cmd("exit 3", time_out=150, sleep_inter=30) == "" and 0.1 < time.time() - t0 < 1.5
    st = StageTimers()
    with st.span("a"):
        time.sleep(0.01)
    st.add("b", 0.5)
    st.add("b", 0.25)
    assert st.count == {"a": 1, "b": 2} and abs(st.total["b"] - 0.75) < 1e-9 and st.total["a"] >= 0.009
    assert "b=750.00ms/2" in st.report()
    ct = CudaStageTimers(True)                       # no GPU here: a disabled timer costs nothing and reports nothing
    with ct.span("x"):
        pass
    assert ct.summary() == {} or "x" in ct.summary()


def test_queue_dataset_stream_can_be_abandoned_and_restarted(tmp_path):
    import torch
    from paddlebox_b200.dataset import DatasetFactory
    from tests.util_data import slot_vars, write_files
    files = write_files(str(tmp_path), n_files=4, lines_per_file=40, n_sparse=2, n_dense=1, key_space=30)
    ds = DatasetFactory().create_dataset("QueueDataset")
    ds.set_use_var(slot_vars(2, 1)); ds.set_batch_size(8); ds.set_thread(1); ds.set_filelist(files)
    for _ in range(3):
        it = ds.batches(device=torch.device("cpu"))
        next(it); next(it)
        it.close()                              # the consumer walks away in the middle of the first group
    assert sum(b.batch_size for b in ds.batches(device=torch.device("cpu"))) == 160
