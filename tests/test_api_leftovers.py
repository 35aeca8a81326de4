"""Public names that no other test or example touched (found by scanning the package for names never mentioned
under tests/, examples/, bench.py): each is exercised once here, so that a typo in a rarely used entry point shows
on the CPU box."""
import io
import os

import numpy as np
import pytest
import torch

from paddlebox_b200.box_wrapper import BoxWrapper
from tests.util_data import slot_vars, write_files


@pytest.fixture
def box():
    BoxWrapper.reset_instance()
    b = BoxWrapper.set_instance(embedx_dim=4, expand_embed_dim=0, device=torch.device("cpu"), rank=0, world_size=1)
    b.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0}, slot_vector=[1, 2, 3], slot_omit_in_feedpass=[3])
    yield b
    BoxWrapper.reset_instance()


def test_wrapper_accessors(box):
    assert (box.get_gpu_num(), box.get_device_count(), box.get_mpi_rank(), box.get_mpi_size()) == (0, 0, 0, 1)
    assert box.get_nccl_rank_id() == 0 and box.get_nccl_rank_id(0) == 0 and box.get_device_id() == 0 == box.get_place_device_id()
    assert box.get_round_id() == 0 and box.get_dataset_id() == 0
    box.begin_feed_pass("20240101"); box.feed_pass(torch.tensor([5, 6])); box.end_feed_pass(); box.begin_pass(); box.end_pass()
    assert box.get_round_id() == 1
    assert box.get_slot_vector() == [1, 2, 3] and box.get_omited_slot() == {3} and box.get_expand_embed_dim() == 0
    assert box.get_feedpass_thread_num() >= 1 and box.gpu_mem_used() == 0
    assert not box.use_afs_api() and box.get_file_mgr() is box.file_mgr and box.use_afs_api()
    assert box.phase_num() == 1 and box.mode() == 0
    box.init_metric("AucCalculator", "j", metric_phase=1)
    box.init_metric("AucCalculator", "u", metric_phase=0)
    assert box.phase_num() == 2
    box.set_test_mode(True)
    assert box.mode() == 1
    box.set_test_mode(False)
    assert torch.equal(box.sync_dense(torch.ones(3)), torch.ones(3))
    got = []
    box.exec_range_func(10, lambda b, e: got.append((b, e)), 2)
    assert sorted(got)[0][0] == 0 and sum(e - b for b, e in got) == 10
    from paddlebox_b200.ops.functional import store_q_value
    store_q_value(torch.tensor([0.5, 0.25]))
    q = box.get_q_tensor()
    assert q is not None


def test_fleet_topology_names_and_table_calls(tmp_path, box):
    from paddlebox_b200 import fleet
    assert fleet.node_num() == 1 and fleet.local_rank() == 0 and fleet.local_device_ids() == [] and fleet.world_device_ids() == [[0]]
    assert fleet.worker_endpoints() and isinstance(fleet.worker_endpoints(True), str)
    assert fleet.server_num() == 0 and fleet.server_index() == 0 and fleet.server_endpoints() == [] and not fleet.is_coordinator()
    m = torch.nn.Linear(2, 2)
    assert fleet.distributed_model(m) is m
    k = np.arange(1, 31, dtype=np.uint64)
    box.host_table.fetch_rows(k, create=True)
    d = str(tmp_path)
    assert fleet.save_one_table(0, d + "/a") == 30 and fleet.save_multi_table_one_path([0], d + "/b") == 30
    assert fleet.save_inference_model(None, d + "/inf") == 30 and fleet.save(d + "/s") == 30
    fleet.clear_one_table(0)
    assert box.host_table.size() == 0 and fleet.load_one_table(0, d + "/a") == 30
    fleet.set_file_num_one_shard(0, 4)
    fleet.run_server(); fleet.init_server(); fleet.init_worker()
    assert fleet.shrink(threshold=1e9) >= 0


def test_role_makers_without_a_launcher(monkeypatch):
    from paddlebox_b200.role_maker import (MPISymetricRoleMaker, PaddleCloudRoleMaker, Role, UserDefinedCollectiveRoleMaker,
                                           UserDefinedRoleMaker)
    u = UserDefinedRoleMaker(current_id=1, role=Role.WORKER, worker_num=3, server_endpoints=["h:1", "h:2"])
    assert (u.is_worker(), u.is_server(), u.worker_index(), u.worker_num(), u.server_num(), u.is_first_worker()) == \
        (True, False, 1, 3, 2, False)
    assert u.role_id() == 1 and "current_id: 1" in u.to_string() and u.rendezvous() == (None, None)
    s = UserDefinedRoleMaker(current_id=1, role=Role.SERVER, worker_num=2, server_endpoints=["h:1", "h:2"])
    assert s.is_server() and s.server_index() == 1 and s.worker_index() == 0
    with pytest.raises(ValueError):
        UserDefinedRoleMaker(current_id=2, role=Role.SERVER, worker_num=2, server_endpoints=["h:1", "h:2"])
    with pytest.raises(TypeError):
        UserDefinedRoleMaker(role=7)
    c = UserDefinedCollectiveRoleMaker(1, ["127.0.0.1:6170", "127.0.0.1:6171"])
    assert c.worker_num() == 2 and c.rendezvous() == ("127.0.0.1", 6170) and c.get_trainer_endpoints()[1].endswith("6171")
    with pytest.raises(ValueError):
        UserDefinedCollectiveRoleMaker(2, ["a:1", "b:2"])
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "3"); monkeypatch.setenv("OMPI_COMM_WORLD_SIZE", "4")
    monkeypatch.setenv("MASTER_ADDR", "10.0.0.1"); monkeypatch.setenv("MASTER_PORT", "1234")
    m = MPISymetricRoleMaker()
    assert (m.worker_index(), m.worker_num(), m.rendezvous()) == (3, 4, ("10.0.0.1", 1234))
    for k in ("PADDLE_PSERVERS_IP_PORT_LIST", "PADDLE_TRAINERS_NUM", "TRAINING_ROLE"):
        monkeypatch.delenv(k, raising=False)
    with pytest.raises(ValueError):
        PaddleCloudRoleMaker().generate_role()
    monkeypatch.setenv("PADDLE_PSERVERS_IP_PORT_LIST", "1.1.1.1:7,1.1.1.2:7"); monkeypatch.setenv("PADDLE_TRAINERS_NUM", "2")
    monkeypatch.setenv("TRAINING_ROLE", "PSERVER"); monkeypatch.setenv("POD_IP", "1.1.1.2"); monkeypatch.setenv("PADDLE_PORT", "7")
    p = PaddleCloudRoleMaker()
    assert p.is_server() and p.server_index() == 1 and p.get_pserver_endpoints() == ["1.1.1.1:7", "1.1.1.2:7"]


def test_data_generators(monkeypatch, capsys):
    from paddlebox_b200.data_generator import DataGenerator, MultiSlotDataGenerator, MultiSlotStringDataGenerator

    class G(MultiSlotDataGenerator):
        def generate_sample(self, line):
            def it():
                a, b = line.split()
                yield [("click", [int(a)]), ("ids", [int(b), int(b) + 1])]
            return it

    g = G()
    g.set_batch(2)
    monkeypatch.setattr("sys.stdin", io.StringIO("1 10\n0 20\n1 30\n"))
    g.run_from_stdin()
    assert capsys.readouterr().out == "1 1 2 10 11\n1 0 2 20 21\n1 1 2 30 31\n"
    out = io.StringIO()
    with pytest.raises(ValueError):
        class Bad(G):
            def generate_sample(self, line):
                return lambda: iter([[("click", [])]])
        Bad().run_from_iterable(["x"], out)

    class S(MultiSlotStringDataGenerator):
        def generate_sample(self, line):
            return lambda: iter([[("w", line.split())]])
    out = io.StringIO()
    S().run_from_iterable(["a b c"], out)
    assert out.getvalue() == "3 a b c\n"
    assert list(DataGenerator().generate_batch([1, 2])()) == [1, 2]
    with pytest.raises(NotImplementedError):
        DataGenerator().generate_sample("x")


def test_dataset_archive_mode_and_setters(tmp_path):
    from paddlebox_b200.dataset import DatasetFactory
    files = write_files(str(tmp_path), n_files=2, lines_per_file=30, n_sparse=2, n_dense=1, key_space=40, ins_id=True)
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var(slot_vars(2, 1)); ds.set_batch_size(8); ds.set_thread(2); ds.set_parse_ins_id(True)
    ds.set_filelist(files)
    ds.set_hdfs_config("hdfs://x", "u,p"); ds.set_download_cmd("cat"); ds.set_input_type(0); ds.set_pipe_command("cat")
    ds.set_archivefile(str(tmp_path / "pass.arch"))
    ds.preload_into_disk()                               # read + parse, spill as a binary archive, free the memory
    assert ds.get_memory_data_size() == 0 and os.path.getsize(tmp_path / "pass.arch") > 0
    ds.wait_load_disk_done()
    assert ds.get_memory_data_size() == 60
    ds.set_merge_by_lineid(2)
    dropped = ds.merge_by_lineid()
    assert dropped >= 0 and ds.get_memory_data_size() == 60 - dropped
    ds.release_memory()
    ds.load_into_disk(str(tmp_path / "p2.arch"))
    assert os.path.getsize(tmp_path / "p2.arch") > 0


def test_graph_store_leftovers(tmp_path):
    from paddlebox_b200.graph import DeviceGraph, GraphDataGenerator, GraphGpuWrapper, GraphTable
    ef = tmp_path / "w.txt"
    ef.write_text("1\t2\t0.5\n1\t3\t2.0\n2\t3\n")
    g = GraphTable()
    assert g.load_weighted_edge_file(str(ef), 0, False) == 3
    g.build(0)
    assert g.is_weighted(0) and g.all_ids(0).tolist() == [1, 2, 3]
    g.add_nodes([1, 2, 3], [[(0, 11)], [(0, 22), (1, 5)], []], ntype=0)
    f = g.node_features(np.array([2, 3, 9], np.uint64), [0, 1], 0)
    assert f[0].tolist() == [22, 5] and f[1].tolist() == [0, 0] and f[2].tolist() == [0, 0]
    g.add_nodes([7, 8], None, ntype=1)
    assert sorted(g.typed_nodes(1).tolist()) == [7, 8]
    dg = g.to_device("cpu", [0])
    assert isinstance(dg, DeviceGraph)
    idx, known = dg.index_of(torch.tensor([3, 1, 99]), 0)
    assert idx.tolist()[:2] == [2, 0] and known.tolist() == [True, True, False]
    w = GraphGpuWrapper()
    w.set_feature_separator(",")
    assert w.feature_separator == ","
    gen = GraphDataGenerator(g, batch_size=2, walk_len=3, window=1, metapath=(0,), device="cpu")
    assert len(gen.start_nodes()) > 0


def test_side_tables_and_tdm_leftovers():
    from paddlebox_b200.ops.tdm import TDMTree
    from paddlebox_b200.side_tables import BoxFileMgr, GpuReplicaCache
    c = GpuReplicaCache(2)
    assert c.add_items([1.0, 2.0]) == 0 and c.add_items([3.0, 4.0]) == 1 and c.size() == 2
    c.to_hbm(torch.device("cpu"))
    assert c.pull_cache_value(torch.tensor([1, 0])).tolist() == [[3.0, 4.0], [1.0, 2.0]]
    c.push_cache_value(torch.tensor([0]), torch.zeros(1, 2))          # a no-op, as in the reference
    m = BoxFileMgr()
    m.init("", "")
    m.destory()
    tree = TDMTree(torch.arange(10, 18), branch=2)
    pos, known = tree.leaf_position(torch.tensor([10, 17, 99]))
    assert known.tolist() == [True, True, False] and pos[0] != pos[1]


def test_transpiler_and_optimizer_names():
    from paddlebox_b200.optimizer import (Adam, BoxPSOptimizer, Collective, DenseOptimizerSpec, GradAllReduce, LocalSGD, MultiThread,
                                          SingleProcessMultiThread)
    spec = Adam(learning_rate=0.01)
    assert isinstance(spec, DenseOptimizerSpec)
    m = torch.nn.Linear(2, 1)
    cfg = BoxPSOptimizer(spec).minimize(m)
    for t, mode in ((GradAllReduce(), "allreduce"), (SingleProcessMultiThread(), "allreduce"), (MultiThread(1, "all_gather"), "allreduce"),
                    (LocalSGD(k_steps=4), "kstep")):
        assert t.transpile(main_program=m) == mode
    assert cfg.sync_weight_step == 4 and cfg.sync_dense_mode == "kstep" and Collective is GradAllReduce
    assert LocalSGD.snapshot_name("w") == "w@SNAPSHOT"


def test_dump_format_flags_and_record_pool_flags(tmp_path, capsys):
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.flags import set_flags
    from paddlebox_b200.utils.dump import FieldDumper
    ids = ["ins1 tagA", "ins2"]
    outs = {"pred.tmp_0": torch.tensor([[0.25], [0.5]]), "label": torch.tensor([[1.0], [0.0]])}

    def dump(flags):
        set_flags(flags)
        d = tmp_path / ("d" + "".join(str(int(v)) for v in flags.values()))
        fd = FieldDumper(str(d), ["pred.tmp_0", "label"], 0, 1, 1, rank=0)
        fd.dump_batch(ids, outs)
        fd.close()
        return sorted(open(os.path.join(d, f)).read() for f in os.listdir(d))[0].splitlines()

    try:
        plain = dump({"dump_filed_same_as_aibox": False, "lineid_have_extend_info": False})
        assert plain == ["ins1 tagA\tpred.tmp_0:1:0.25\tlabel:1:1", "ins2\tpred.tmp_0:1:0.5\tlabel:1:0"]
        aibox = dump({"dump_filed_same_as_aibox": True, "lineid_have_extend_info": True})
        assert aibox == ["ins1\tpred:0.25\tlabel:1\ttagA", "ins2\tpred:0.5\tlabel:0"]
        set_flags({"enable_print_dump_info_debug": True, "enable_print_dump_field_debug": True})
        dump({"dump_filed_same_as_aibox": False, "lineid_have_extend_info": False})
        out = capsys.readouterr().out
        assert "dump ins count=2 of 2" in out and "name=label width=1" in out
    finally:
        set_flags({"dump_filed_same_as_aibox": False, "lineid_have_extend_info": False, "enable_print_dump_info_debug": False,
                   "enable_print_dump_field_debug": False})
    # record pool: released records are kept for the next pass, unless the recycle flags say otherwise
    files = write_files(str(tmp_path / "data") if (tmp_path / "data").mkdir() is None else "", n_files=1, lines_per_file=50,
                        n_sparse=2, n_dense=1, key_space=20)
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_use_var(slot_vars(2, 1)); ds.set_batch_size(8); ds.set_thread(1); ds.set_filelist(files)
    ds.load_into_memory(); ds.release_memory()
    assert ds._native.idle_records() == 50
    try:
        set_flags({"enable_force_mem_recyle": True})
        ds.load_into_memory(); ds.release_memory()
        assert ds._native.idle_records() == 0
        set_flags({"enable_force_mem_recyle": False, "padbox_record_pool_max_size": 10})
        ds.load_into_memory(); ds.release_memory()
        assert ds._native.idle_records() == 10
    finally:
        set_flags({"enable_force_mem_recyle": False, "padbox_record_pool_max_size": 2000000})


def test_trainer_desc_text_and_its_fields(tmp_path):
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.ops.pull_box_sparse import SlotBatch
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    text = """
      class_name: "BoxPSTrainer"  device_worker_name: "BoxPSWorker"  thread_num: 8  debug: true
      dump_fields_path: "%s"  dump_fields: "pred"  dump_fields: "label"  dump_param: "fc_0.w_0"
      user_define_dump_filename: "join"  enable_random_dump: true  random_with_lineid: true  dump_interval: 3
      check_nan_var_names: "pred"  loss_names: "loss"  scale_sparse_gradient_with_batch_size: false
      boxps_param { async_mode: false sync_dense_mode: 1 sync_weight_step: 4 dump_thread_num: 2 param_need_sync: "fc_0.w_0" }
    """ % str(tmp_path / "dump")
    cfg = TrainerConfig.from_desc(text)
    assert cfg.debug and cfg.dump_fields == ["pred", "label"] and cfg.dump_param == ["fc_0.w_0"] and cfg.dump_mode == 1
    assert cfg.dump_interval == 3 and cfg.user_define_dump_filename == "join" and cfg.check_nan_var_names == ["pred"]
    assert (cfg.sync_dense_mode, cfg.sync_weight_step, cfg.dump_thread_num) == ("kstep", 4, 2)
    assert cfg.scale_sparse_gradient_with_batch_size is False and cfg.loss_names == ["loss"]
    assert TrainerConfig.from_desc('boxps_param { async_mode: true }').sync_dense_mode == "async"
    assert TrainerConfig.from_desc('debug: false').dump_mode == 0

    def one_step(scale_with_batch):
        BoxWrapper.reset_instance()
        box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
        box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0, "initial_range": 0.0, "mf_initial_range": 0.01})
        keys = torch.tensor([3, 4, 5, 6], dtype=torch.int64)
        batch = SlotBatch(keys, torch.tensor([0, 1, 2, 3, 4], dtype=torch.int32), 2, 2, torch.tensor([[1.0, .5], [0.0, .25]]),
                          ins_ids=["a x", "b"])
        box.begin_feed_pass("20240101"); box.feed_pass(keys); box.end_feed_pass(); box.begin_pass()
        torch.manual_seed(0)
        c = TrainerConfig(dump_fields=["pred"], dump_fields_path=str(tmp_path / f"d{int(scale_with_batch)}"),
                          user_define_dump_filename="join", scale_sparse_gradient_with_batch_size=scale_with_batch)
        tr = CTRTrainer(CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(4,)), box, c)
        before = box.host_table.fetch_rows(keys.numpy().astype(np.uint64), create=False).copy()
        tr.train_step(batch)
        tr.train_step(batch)                                   # (the first push of a key only creates its state)
        if tr.dumper is not None:
            tr.dumper.close()
        box.end_pass()
        after = box.host_table.fetch_rows(keys.numpy().astype(np.uint64), create=False)
        BoxWrapper.reset_instance()
        return np.abs(after[:, 2] - before[:, 2])            # movement of embed_w

    big, small = one_step(True), one_step(False)
    assert np.all(big > 0) and np.all(small > 0) and np.all(big > small)          # -batch_size vs -1 as the push scale
    assert os.listdir(tmp_path / "d1") == ["part-join-00000"]


def test_psgpu_begin_pass_without_a_queued_task_is_an_error_not_a_hang():
    from paddlebox_b200.ps_gpu import PSGPUWrapper
    PSGPUWrapper.reset_instance()
    try:
        ps = PSGPUWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"), rank=0, world_size=1)
        ps.init_gpu_ps(conf={})
        ps.build_task(np.array([1, 2, 3], np.uint64))
        ps.end_pass()
        with pytest.raises(RuntimeError):
            ps.begin_pass()                    # nothing was queued for a second pass
        ps.finalize()
    finally:
        PSGPUWrapper.reset_instance()


def test_dataset_generator_check_and_local_tables(tmp_path):
    from paddlebox_b200.data_generator import MultiSlotDataGenerator
    from paddlebox_b200.dataset import DatasetFactory, dense_slot, sparse_slot

    class Gen(MultiSlotDataGenerator):
        def generate_sample(self, line):
            def it():
                a, b = line.split()
                yield [("ids", [int(a)]), ("score", [float(b)])]
            return it

    class BadGen(Gen):
        def generate_sample(self, line):
            def it():
                yield [("ids", [1]), ("score", [2])]               # an int in a float slot
            return it
    f = tmp_path / "part"
    f.write_text("3 0.5\n4 1.5\n")
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    use = [sparse_slot("ids"), dense_slot("score", 1)]
    ds._check_use_var_with_data_generator(use, Gen(), str(f))
    with pytest.raises(TypeError):
        ds._check_use_var_with_data_generator(use, BadGen(), str(f))
    with pytest.raises(ValueError):
        ds._check_use_var_with_data_generator(use[:1], Gen(), str(f))
    ds.set_use_var(use)
    assert ds._desc() == ds.desc()


def test_boxps_helper_object_drives_the_pass_lifecycle(tmp_path):
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import BoxPS, DatasetFactory
    from tests.util_data import slot_vars, write_files
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"))
    box.initialize_gpu_and_load_model({})
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var(slot_vars(3, 1))
    ds.set_batch_size(8)
    ds.set_filelist(write_files(str(tmp_path), n_files=1, lines_per_file=40, n_sparse=3, n_dense=1, key_space=30, seed=1))
    helper = BoxPS(ds)
    helper.set_date(2024, 1, 2)
    assert ds._date == "20240102"
    helper.preload_into_memory()
    helper.wait_feed_pass_done()
    helper.begin_pass()
    assert ds.get_memory_data_size() == 40 and box.host_table.size() > 0
    helper.end_pass(False)
    with pytest.raises(TypeError):
        BoxPS(DatasetFactory().create_dataset("QueueDataset"))
    BoxWrapper.reset_instance()


def test_core_metric_registry_and_dataset_getters(tmp_path):
    import torch
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.dataset import DatasetFactory
    from paddlebox_b200.metrics import Metric
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig
    from tests.util_data import slot_vars, write_files
    Metric.reset_instance()
    m = Metric()
    assert Metric() is m and Metric.get_instance() is m
    m.init_metric("AucCalculator", "auc", "label", "pred", "", "", "", 1, "", False, 100000)
    m.init_metric("WuAucCalculator", "wu", "label", "pred", "", "", "uid", 0, "", False)
    assert m.get_metric_name_list(1) == ["auc"] and m.get_metric_name_list(-1) == ["auc", "wu"]
    with pytest.raises(ValueError):
        m.get_metric_msg("nosuch")
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"))
    box.initialize_gpu_and_load_model({})
    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var(slot_vars(3, 1))
    ds.set_batch_size(16)
    ds.set_thread(2)
    files = write_files(str(tmp_path), n_files=1, lines_per_file=64, n_sparse=3, n_dense=1, key_space=30, seed=1)
    ds.set_filelist(files)
    ds.set_hdfs_config("afs://x", "u,p")
    assert ds.get_filelist() == files and ds.get_thread_num() == 2 and ds.get_trainer_num() == 1
    assert ds.get_hdfs_config() == ("afs://x", "u,p") and ds.get_download_cmd() == ds.download_cmd
    assert ds.get_data_feed_desc() == ds.desc() and ds.get_fleet_send_batch_size() > 0
    ds.register_client2client_msg_handler()
    ds.load_into_memory()
    ds.begin_pass()
    CTRTrainer(CtrDnn(num_slots=3, embedx_dim=4, dense_dim=1, hidden=(4,)), box, TrainerConfig()).train_from_dataset(ds)
    ds.end_pass(False)
    msg = m.get_metric_msg("auc")                               # the trainer fed the registry as well
    assert len(msg) == 8 and msg[7] == 64 and 0.0 <= msg[0] <= 1.0
    assert m.get_metric_msg("auc")[7] == 0                       # reading resets
    m.flip_phase()
    assert m.phase() == 0
    Metric.reset_instance()
    BoxWrapper.reset_instance()
