"""FleetUtil / GPUPSUtil: the job-script layer around the wrapper (reference fleet_util.py)."""
import json
import os

import numpy as np
import pytest
import torch

from paddlebox_b200 import fleet_util as FU
from paddlebox_b200.box_wrapper import BoxWrapper


def test_brace_expansion_and_online_pass_interval():
    assert FU.expand_braces("{20190720..20190722}") == ["20190720", "20190721", "20190722"]
    assert FU.expand_braces("{0..3}") == ["0", "1", "2", "3"] and FU.expand_braces("{08..10}") == ["08", "09", "10"]
    assert FU.expand_braces("{a,b}") == ["a", "b"] and FU.expand_braces("7") == ["7"]
    u = FU.FleetUtil()
    # the reference's docstring example: 5-minute splits, 2 per pass -> 144 passes of 2 splits
    iv = u.get_online_pass_interval("{20190720..20190729}", "{0..23}", 5, 2, False)
    assert len(iv) == 144 and iv[0] == ["0000", "0005"] and iv[1] == ["0010", "0015"] and iv[-1] == ["2350", "2355"]
    hourly = u.get_online_pass_interval("{20190720..20190720}", "{0..23}", 60, 1, True)
    assert hourly == [["%02d" % h] for h in range(24)]
    part = u.get_online_pass_interval("x", "{8..9}", 30, 2, False)
    assert part == [["0800", "0830"], ["0900", "0930"]]


def test_metrics_from_raw_statistics_match_the_host_calculator():
    from paddlebox_b200.host import HostAuc
    rng = np.random.default_rng(0)
    n, nb = 20000, 1000
    p = rng.random(n).astype(np.float32) ** 2
    y = (rng.random(n) < p).astype(np.float32)
    calc = HostAuc(nb)
    calc.add(p, y)
    want = calc.compute()                      # [auc, bucket_error, mae, rmse, actual, predicted, copc, size]
    idx = np.minimum((p.astype(np.float64) * nb).astype(np.int64), nb - 1)
    pos = np.bincount(idx, weights=y, minlength=nb)
    neg = np.bincount(idx, weights=1 - y, minlength=nb)
    d = p.astype(np.float64) - y
    got = FU.FleetUtil.metrics_from_stats(pos, neg, float((d * d).sum()), float(np.abs(d).sum()), float(p.sum()),
                                          float(p.sum()), float(y.sum()), float(n))
    assert got[8] == n and abs(got[0] - want[0]) < 1e-9 and abs(got[1] - want[1]) < 1e-9
    assert np.allclose(got[2:7], want[2:7], rtol=1e-5)
    assert FU.auc_from_buckets(np.zeros(8), np.ones(8)) == 0.5          # one class absent
    assert FU.FleetUtil.metrics_from_stats(pos * 0, neg * 0, 0, 0, 0, 0, 0, 0)[0] == 0.5


@pytest.fixture
def box():
    BoxWrapper.reset_instance()
    b = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    b.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
    yield b
    BoxWrapper.reset_instance()


def _touch(box, keys, clicks=1.0):
    k = np.asarray(keys, dtype=np.uint64)
    box.host_table.fetch_rows(k, create=True)
    push = np.zeros((k.size, box.host_table.layout.push_stride), np.float32)
    push[:, 1] = 3.0
    push[:, 2] = clicks
    push[:, 3:] = 0.05
    box.host_table.push(k, torch.from_numpy(push))


def test_a_training_day_through_fleet_util(tmp_path, box):
    out = str(tmp_path / "output")
    u = FU.FleetUtil(box=box)
    assert u.get_last_save_model(out)[:3] == [-1, -1, ""] and u.get_last_save_xbox_base(out)[:2] == [-1, -1]
    key = 1563000000
    _touch(box, range(1, 101))
    # pass 1: delta + patch donefile
    assert u.save_delta_model(out, 20190722, 1) == 100
    assert os.path.isdir(f"{out}/20190722/delta-1")
    assert u.write_xbox_donefile(out, 20190722, 1, key, ["/data/a", "/data/b"], hadoop_fs_name="hdfs://ns")
    assert not u.write_xbox_donefile(out, 20190722, 1, key, "/data/a")          # not newer: refused
    _touch(box, range(50, 121))
    assert u.save_delta_model(out, 20190722, 2) == 71                            # only what changed since pass 1
    assert u.write_xbox_donefile(out, 20190722, 2, key, "/data/c", hadoop_fs_name="hdfs://ns")
    lines = open(f"{out}/xbox_patch_done.txt").read().strip().split("\n")
    assert len(lines) == 2
    e = json.loads(lines[0])
    assert list(e) == ["id", "key", "input", "record_count", "partition_type", "job_name", "ins_tag", "ins_path",
                       "job_id", "monitor_data", "monitor_path", "mpi_size"]
    assert e["input"] == f"hdfs://ns{out}/20190722/delta-1/000" and e["key"] == str(key) and e["ins_path"] == "/data/a,/data/b"
    assert e["record_count"] == "111111" and e["partition_type"] == "2" and e["mpi_size"] == "1"
    assert u.get_last_save_xbox(out) == [20190722, 2, f"hdfs://ns{out}/20190722/delta-2", key]
    # per-pass full model + donefile.txt
    assert u.save_model(out, 20190722, 2) == 120
    assert u.write_model_donefile(out, 20190722, 2, key)
    assert not u.write_model_donefile(out, 20190722, 2, key)
    # end of day: base + batch model
    assert u.save_xbox_base_model(out, 20190722) == 120 and os.path.isdir(f"{out}/20190722/base")
    assert u.write_xbox_donefile(out, 20190722, -1, key, "/data")
    base = json.loads(open(f"{out}/xbox_base_done.txt").read().strip())
    assert base["id"] == str(key) and base["input"].endswith("/20190722/base/000")
    assert u.get_last_save_xbox_base(out) == [20190722, f"{out}/20190722/base", key]
    assert u.save_batch_model(out, 20190722) == 120 and os.path.isdir(f"{out}/20190722/0")
    assert u.write_model_donefile(out, 20190722, -1, key)
    rows = [r.split("\t") for r in open(f"{out}/donefile.txt").read().strip().split("\n")]
    assert rows == [["20190722", str(key), f"{out}/20190722/2/", "2", "0"],
                    ["20190722", str(key), f"{out}/20190722/0/", "-1", "0"]]
    assert u.get_last_save_model(out) == [20190722, -1, f"{out}/20190722/0/", key]
    # cache model: hot rows only, and its meta file
    _touch(box, range(1, 11), clicks=40.0)
    n_hot = u.save_cache_model(out, 20190722, 2, cache_threshold=30.0)
    assert n_hot == 10
    assert u.write_cache_donefile(out, 20190722, 2, n_hot) and not u.write_cache_donefile(out, 20190722, 2, n_hot)
    assert open(f"{out}/20190722/delta-2/000_cache/sparse_cache.meta").read() == "file_prefix:part\npart_num:16\nkey_num:10\n"
    # restart: a fresh wrapper resumes from what donefile.txt names
    want = box.host_table.fetch_rows(np.arange(1, 121, dtype=np.uint64), create=False)
    BoxWrapper.reset_instance()
    b2 = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
    b2.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
    u2 = FU.GPUPSUtil(box=b2)
    day, pass_id, path, k = u2.get_last_save_model(out)
    assert (day, pass_id, k) == (20190722, -1, key)
    assert u2.load_model(out, day, pass_id) == 120
    got = b2.host_table.fetch_rows(np.arange(1, 121, dtype=np.uint64), create=False)
    # (the batch model was saved before the last 10 hot pushes)
    assert np.allclose(got[10:], want[10:], rtol=1e-6)


def test_global_metrics_come_from_the_metric_registry(box, capsys):
    box.init_metric("AucCalculator", "join", "label", "pred", metric_phase=1, bucket_size=10000)
    rng = np.random.default_rng(1)
    p = torch.from_numpy(rng.random(5000).astype(np.float32))
    y = (torch.from_numpy(rng.random(5000).astype(np.float32)) < p).float()
    box.get_metric("join").add_data({"pred": p, "label": y})
    u = FU.FleetUtil(box=box)
    m = u.print_global_metrics("join", print_prefix="pass 1")
    out = capsys.readouterr().out
    assert "pass 1 global AUC=" in out and "Ins number=5000" in out
    assert len(m) == 9 and 0.7 < m[0] < 0.9 and m[8] == 5000 and abs(m[4] - float(y.mean())) < 1e-6
    assert u.get_global_metrics("nope") == [None] * 9
    box.get_metric("join").add_data({"pred": p, "label": y})
    u.set_zero("join")
    assert u.get_global_metrics("join")[8] == 0


def test_full_models_carry_the_dense_parameters_of_the_trainer(tmp_path, box):
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    out = str(tmp_path / "o")
    torch.manual_seed(0)
    tr = CTRTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=2, hidden=(8, 4)), box, TrainerConfig())
    _touch(box, range(1, 20))
    u = FU.FleetUtil(box=box, trainer=tr)
    u.save_model(out, 20240101, 3)
    assert os.path.isdir(f"{out}/20240101/3/dense")
    assert set(u.pull_all_dense_params()) == {n for n, _ in tr.model.named_dense_params()}
    plug = u.save_paddle_params(out, 20240101, 3)
    assert plug.endswith("/20240101/3/dnn_plugin") and os.listdir(plug)
    want = tr.store.flat.clone()
    torch.manual_seed(1)
    tr2 = CTRTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=2, hidden=(8, 4)), box, TrainerConfig())
    assert not torch.equal(tr2.store.flat, want)
    FU.FleetUtil(box=box, trainer=tr2).load_model(out, 20240101, 3)
    assert torch.allclose(tr2.store.flat, want)


def test_pslib_style_fleet_calls_act_on_the_wrapper(tmp_path, box):
    from paddlebox_b200 import fleet
    from paddlebox_b200.ops.functional import DataNorm
    _touch(box, range(1, 51))
    d0 = str(tmp_path / "full")
    assert fleet.save_persistables(None, d0, mode=0) == 50
    assert fleet.save_persistables(None, str(tmp_path / "delta"), mode=1) == 50
    assert fleet.save_persistables(None, str(tmp_path / "delta2"), mode=1) == 0      # the baseline moved
    _touch(box, range(1, 6), clicks=50.0)
    assert fleet.save_cache_model(None, str(tmp_path / "cache"), cache_threshold=30.0) == 5
    st = fleet.print_table_stat(0)
    assert st["rows"] == 50 and st["rank"] == 0
    fleet.clear_model()
    assert box.host_table.size() == 0
    assert fleet.load_model(d0) == 50 and box.host_table.size() == 50
    removed = fleet.shrink_sparse_table(decay=0.5, delete_threshold=1e9)              # nothing survives this bar
    assert removed == 50 and box.host_table.size() == 0
    fleet.set_date(0, 20240105)

    class T:                                                                          # a trainer-shaped holder
        model = torch.nn.Sequential(DataNorm(4))
    before = T.model[0].batch_size.clone()
    fleet.shrink_dense_table(0.9, trainer=T)
    assert torch.allclose(T.model[0].batch_size, before * 0.9)
    from paddlebox_b200.optimizer import BoxPSOptimizer
    assert isinstance(fleet.distributed_optimizer(object()), BoxPSOptimizer)


FLEET_DESC = """
# what a pslib job hands its trainers (server part only)
server_param {
  downpour_server_param {
    downpour_table_param {
      table_id: 0
      table_class: "DownpourSparseTable"
      shard_num: 1950
      type: PS_SPARSE_TABLE
      accessor {
        accessor_class: "DownpourUnitAccessor"
        fea_dim: 11
        embedx_dim: 8
        embedx_threshold: 5
        downpour_accessor_param {
          nonclk_coeff: 0.2
          click_coeff: 2.0
          base_threshold: 1.5
          delta_threshold: 0.25
          delta_keep_days: 16.0
          show_click_decay_rate: 0.98
          delete_threshold: 0.8
          delete_after_unseen_days: 30
        }
        embed_sgd_param {
          name: "adagrad"
          adagrad { learning_rate: 0.05 initial_g2sum: 3.0 initial_range: 0.0001 weight_bounds: -10.0 weight_bounds: 10.0 }
        }
        embedx_sgd_param {
          name: "adam"
          adam { learning_rate: 0.001, initial_range: 1e-4; beta1_decay_rate: 0.9 beta2_decay_rate: 0.999 ada_epsilon: 1e-8
                 weight_bounds: [-5.0, 5.0] }
        }
      }
    }
    downpour_table_param { table_id: 1 table_class: "DownpourDenseTable" type: PS_DENSE_TABLE
      accessor { accessor_class: "DownpourDenseValueAccessor" dense_sgd_param { name: "adam" } fea_dim: 1234 } }
    service_param { server_class: "DownpourBrpcPsServer" start_server_port: 0 }
  }
}
trainer_param { dense_table { table_id: 1 dense_variable_name: "fc_0.w_0" dense_variable_name: "fc_0.b_0" } }
"""


def test_fleet_desc_text_gives_the_sparse_config():
    from paddlebox_b200.fleet_desc import parse_text_format, sparse_config_from_fleet_desc
    d = parse_text_format(FLEET_DESC)
    tabs = d["server_param"]["downpour_server_param"]["downpour_table_param"]
    assert len(tabs) == 2 and tabs[0]["type"] == "PS_SPARSE_TABLE" and tabs[0]["shard_num"] == 1950
    assert d["trainer_param"]["dense_table"]["dense_variable_name"] == ["fc_0.w_0", "fc_0.b_0"]
    conf, opt, extra = sparse_config_from_fleet_desc(FLEET_DESC)
    assert opt == "adagrad" and extra == {"embedx_dim": 8}
    assert conf == {"nonclk_coeff": 0.2, "clk_coeff": 2.0, "mf_create_thresholds": 5.0, "learning_rate": 0.05,
                    "initial_range": 0.0001, "initial_g2sum": 3.0, "min_bound": -10.0, "max_bound": 10.0,
                    "mf_learning_rate": 0.001, "mf_initial_range": 1e-4, "mf_beta1_decay_rate": 0.9,
                    "mf_beta2_decay_rate": 0.999, "mf_ada_epsilon": 1e-8, "mf_min_bound": -5.0, "mf_max_bound": 5.0}
    # the CTR accessors use one rule for embed and embedx
    ctr = FLEET_DESC.replace("DownpourUnitAccessor", "DownpourCtrAccessor").replace(
        "embed_sgd_param {", "sparse_sgd_param { learning_rate: 0.07 initial_g2sum: 2.0 initial_range: 0.001 "
                             "weight_bounds: -1.0 weight_bounds: 1.0 }\n embed_sgd_param {")
    c2, o2, _ = sparse_config_from_fleet_desc(ctr)
    assert o2 == "adagrad" and c2["learning_rate"] == c2["mf_learning_rate"] == 0.07 and c2["mf_max_bound"] == 1.0
    import pytest
    with pytest.raises(ValueError):
        sparse_config_from_fleet_desc("server_param { }")
    with pytest.raises(ValueError):
        parse_text_format("a { b: 1 ")
    # and the wrapper takes the text as it is
    from paddlebox_b200.ps_gpu import PSGPUWrapper
    PSGPUWrapper.reset_instance()
    try:
        ps = PSGPUWrapper.set_instance(embedx_dim=8, device=torch.device("cpu"), rank=0, world_size=1)
        ps.init_gpu_server(FLEET_DESC)
        cfg = ps.host_table.config
        assert abs(cfg.clk_coeff - 2.0) < 1e-6 and abs(cfg.mf_learning_rate - 0.001) < 1e-9 and abs(cfg.mf_create_thresholds - 5.0) < 1e-6
    finally:
        PSGPUWrapper.reset_instance()


def test_pslib_strategy_configures_the_table_and_the_day_end_thresholds(box):
    from paddlebox_b200 import fleet
    from paddlebox_b200.fleet_desc import sparse_config_from_strategy
    conf, kind, extra, policy = sparse_config_from_strategy({})
    assert kind == "adagrad" and extra == {"embedx_dim": 8} and conf["learning_rate"] == 0.05 and conf["mf_create_thresholds"] == 10
    assert policy["delete_threshold"] == 0.8 and policy["base_threshold"] == 1.5 and policy["show_click_decay_rate"] == 0.98
    c2, k2, _, _ = sparse_config_from_strategy({"sparse_accessor_class": "DownpourUnitAccessor", "sparse_optimizer": "adagrad",
                                               "embedx_sparse_optimizer": "adam", "embedx_sparse_learning_rate": 0.002,
                                               "sparse_weight_bounds": [-2, 2]})
    assert k2 == "adagrad" and c2["mf_learning_rate"] == 0.002 and c2["mf_beta2_decay_rate"] == 0.999 and c2["max_bound"] == 2.0
    try:
        fleet.distributed_optimizer(object(), {"sparse_learning_rate": 0.2, "sparse_click_coeff": 3, "sparse_delete_threshold": 2.5,
                                               "sparse_show_click_decay_rate": 0.5, "sparse_base_threshold": 30.0})
        assert abs(box.host_table.config.learning_rate - 0.2) < 1e-7 and abs(box.host_table.config.clk_coeff - 3.0) < 1e-7
        _touch(box, range(1, 11), clicks=1.0)            # score = 0.1 * (3 - 1) + 3 * 1 = 3.2
        _touch(box, range(11, 21), clicks=0.0)           # score = 0.1 * 3 = 0.3
        # serving base keeps rows over the strategy's base threshold only (none reaches 30)
        import tempfile
        with tempfile.TemporaryDirectory() as d:
            assert fleet.save_persistables(None, d + "/base", mode=2) == 0
            assert fleet.save_persistables(None, d + "/base2", mode=2, base_threshold=1.0) == 10
        # shrink with the strategy's decay 0.5 and delete threshold 2.5: scores become 1.6 and 0.15 -> all removed
        assert fleet.shrink_sparse_table() == 20
    finally:
        fleet.table_policy.clear()


def _free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("kind", ["general", "collective"])
def test_role_makers_join_the_world_from_the_paddle_launcher_environment(kind, tmp_path):
    """Two trainer processes (and, for the PS-style role maker, one process launched as a parameter server) started
    with the variables Paddle's launchers export; ``fleet.init(role_maker)`` joins them into one world."""
    import subprocess
    import sys
    import json as _json
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    worker = os.path.join(root, "tests", "rm_worker.py")
    eps = f"127.0.0.1:{_free_port()},127.0.0.1:{_free_port()}"
    base = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    base.update({"PADDLE_TRAINER_ENDPOINTS": eps, "PADDLE_TRAINERS_NUM": "2", "PADDLE_PSERVERS_IP_PORT_LIST": "127.0.0.1:7001"})
    procs = []
    for r in range(2):
        env = dict(base, PADDLE_TRAINER_ID=str(r), TRAINING_ROLE="TRAINER", PADDLE_TRAINING_ROLE="TRAINER")
        procs.append(subprocess.Popen([sys.executable, str(worker), kind], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    if kind == "general":
        env = dict(base, TRAINING_ROLE="PSERVER", POD_IP="127.0.0.1", PADDLE_PORT="7001")
        procs.append(subprocess.Popen([sys.executable, str(worker), kind], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        so, se = p.communicate(timeout=300)
        assert p.returncode == 0, se[-2000:]
        outs.append(_json.loads(so.strip().splitlines()[-1]))
    for r in range(2):
        o = outs[r]
        assert (o["rank"], o["world"], o["first"]) == (r, 2, r == 0) and o["sum"] == [3.0, 20.0] and o["gathered"] == [0, 1]
        assert o["eps"] == eps.split(",")
    if kind == "general":
        assert outs[2] == {"server": 0}


def test_predictor_follows_a_job_through_its_donefiles(tmp_path):
    """Serving side of the loop: the online example trains two days; a predictor built from the donefiles loads
    the last base; a further day's deltas are picked up by ``refresh_from_donefiles``."""
    import subprocess
    import sys
    from paddlebox_b200.models.ctr import build_model
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch
    from paddlebox_b200.serving import Predictor, serving_models_from_donefiles
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    out = str(tmp_path / "job")
    cmd = [sys.executable, os.path.join(root, "examples", "online_train.py"), "--cpu", "--out", out, "--lines", "128"]
    assert subprocess.run(cmd + ["--days", "20240101,20240102"], capture_output=True, text=True, env=env, timeout=600).returncode == 0
    base, deltas, dense = serving_models_from_donefiles(out)
    assert base == f"{out}/20240102/base" and deltas == [] and dense == f"{out}/20240102/0/dense"
    BoxWrapper.reset_instance()
    model = build_model("deepfm", num_slots=6, embedx_dim=8, dense_dim=3, hidden=(64, 32))
    p = Predictor.from_donefiles(model, out, embedx_dim=8, device=torch.device("cpu"))
    try:
        n0 = p.num_features
        assert n0 > 100 and p.refresh_from_donefiles() == 0
        keys = torch.tensor(p.box.host_table.keys()[:12].astype(np.int64))
        batch = SlotBatch(keys, torch.arange(0, 13, dtype=torch.int32), 2, 6, torch.rand(2, 4))
        before = p.predict(batch).clone()
        assert before.shape == (2,) and torch.all((before > 0) & (before < 1))
        # the job trains one more day: two new deltas appear in xbox_patch_done.txt (and a new base at its end)
        assert subprocess.run(cmd + ["--days", "20240101,20240102,20240103"], capture_output=True, text=True, env=env,
                              timeout=600).returncode == 0
        # (serving_models_from_donefiles now names the new base; the live predictor follows the deltas of its own base)
        p._donefile_root = out
        lines = [ln for ln in open(f"{out}/xbox_patch_done.txt").read().split("\n") if ln.strip()]
        assert len(lines) == 6
        import json as _json
        new_deltas = [os.path.dirname(_json.loads(ln)["input"]) for ln in lines[-2:]]
        got = sum(p.apply_delta(d) > 0 for d in new_deltas)
        assert got == 2 and p.num_features >= n0
        after = p.predict(batch)
        assert not torch.equal(before, after)                    # the rows moved with the new day's training
    finally:
        p.close()


def test_predictor_refresh_picks_up_new_deltas(tmp_path, box):
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.serving import Predictor
    out = str(tmp_path / "o")
    u = FU.FleetUtil(box=box)
    _touch(box, range(1, 41))
    u.save_xbox_base_model(out, 20240101); u.write_xbox_donefile(out, 20240101, -1, 7, "/d")
    _touch(box, range(30, 61))
    u.save_delta_model(out, 20240102, 1); u.write_xbox_donefile(out, 20240102, 1, 7, "/d")
    BoxWrapper.reset_instance()
    p = Predictor.from_donefiles(CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(4,)), out, embedx_dim=4,
                                 device=torch.device("cpu"))
    try:
        assert p.num_features == 61 and p.refresh_from_donefiles() == 0        # 60 keys + the always-present key 0
        # the training job (another process in real life) publishes the next delta
        b2 = BoxWrapper(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
        b2.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0})
        u2 = FU.FleetUtil(box=b2)
        _touch(b2, range(55, 81))
        u2.save_delta_model(out, 20240102, 2); u2.write_xbox_donefile(out, 20240102, 2, 7, "/d")
        assert p.refresh_from_donefiles() == 1 and p.num_features == 81 and p.refresh_from_donefiles() == 0
    finally:
        p.close()


def test_fleet_whitelist_save_split_files_and_small_calls(tmp_path):
    import numpy as np
    from paddlebox_b200 import fleet
    from paddlebox_b200.flags import FLAGS
    from paddlebox_b200.host import HostTable
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"))
    box.initialize_gpu_and_load_model({})
    keys = np.arange(1, 41, dtype=np.uint64)
    rows = box.host_table.fetch_rows(keys, create=True)
    rows[:, 0] = 5.0
    box.host_table.store_rows(keys, rows)
    wl = tmp_path / "white.txt"
    wl.write_text("".join(f"{k}\t1\n" for k in (3, 7, 11, 1000)))
    assert fleet.load_pslib_whitelist(0, str(wl)) == 4
    n = fleet.save_model_with_whitelist(None, str(tmp_path / "wl_model"), mode=0)
    assert n == 3                                                          # key 1000 is not in the table
    back = HostTable(box.host_table.embedx_dim, box.host_table.optimizer, box.host_table.config)
    assert back.load(str(tmp_path / "wl_model")) == 3 and sorted(back.keys().tolist()) == [3, 7, 11]
    np.testing.assert_allclose(back.fetch_rows(np.array([7], np.uint64), create=False), rows[6:7])
    assert fleet.save_model_with_whitelist(None, str(tmp_path / "wl2"), str(wl), mode=0) == 3
    assert box.host_table.size() == 40                                     # the table itself is untouched

    assert fleet.split_files(["a", "b", "c"]) == ["a", "b", "c"]
    with pytest.raises(TypeError):
        fleet.split_files("a,b")
    out = np.zeros(2)
    fleet.all_reduce_worker(np.array([1.0, 2.0]), out)
    assert out.tolist() == [1.0, 2.0] and fleet.is_xpu() is False
    old = FLAGS.pbx_pass_fetch_threads
    fleet.set_pull_local_thread_num(3)
    assert FLAGS.pbx_pass_fetch_threads == 3
    FLAGS.pbx_pass_fetch_threads = old
    fleet._set_client_communication_config(1000, 2000, 3)
    fleet.confirm()
    fleet.revert()
    env = fleet.get_dist_env()
    assert env["trainer_id"] == 0 and env["num_trainers"] == 1
    opt = fleet.distributed_optimizer(__import__("paddlebox_b200.optimizer", fromlist=["x"]).DenseOptimizerSpec("adam", 1e-3))
    m = torch.nn.Linear(2, 1)
    cfg = fleet.minimize(m)
    assert m._pipeline_opt is cfg and opt is fleet._last_optimizer
    assert abs(fleet.metrics.mae([2.0], [4.0]) - 0.5) < 1e-12
    BoxWrapper.reset_instance()
