"""TrainerDesc / DeviceWorker / TrainerFactory and DataFeedDesc: the description objects of the reference's executor."""
import pytest
import torch

from paddlebox_b200.data_feed_desc import DataFeedDesc
from paddlebox_b200.fleet_desc import format_text, parse_text_format
from paddlebox_b200.trainer import TrainerConfig
from paddlebox_b200.trainer_desc import (BoxPSTrainer, BoxPSWorker, DeviceWorkerFactory, DownpourSGD, Hogwild, MultiTrainer,
                                         TrainerFactory)


def test_format_text_round_trips_through_the_parser():
    d = {"name": 'a "quoted" \\ name', "n": 3, "f": 0.5, "flag": True, "rep": [1, 2, 3], "sub": {"x": "y", "deep": {"k": False}},
         "msgs": [{"a": 1}, {"a": 2}], "skip": None}
    back = parse_text_format(format_text(d))
    d.pop("skip")
    assert back == d


def test_factory_defaults_and_opt_info():
    t = TrainerFactory()._create_trainer()
    assert isinstance(t, MultiTrainer) and isinstance(t._device_worker, Hogwild)
    t._set_infer(True)
    t._gen_trainer_desc()
    d = parse_text_format(t._desc())
    assert d["class_name"] == "MultiTrainer" and d["device_worker_name"] == "HogwildWorker" and "push_sparse" in d["hogwild_param"]["skip_ops"]

    model = torch.nn.Linear(2, 1)
    model._pipeline_opt = {"sync_dense_mode": 1, "sync_weight_step": 4, "async_mode": False, "dump_thread_num": 3}
    opt_info = {"trainer": "BoxPSTrainer", "device_worker": "BoxPSWorker", "dump_fields": ["pred", "label"], "dump_fields_path": "/tmp/d",
                "dump_param": ["fc.w"], "enable_random_dump": True, "random_with_lineid": True, "dump_interval": 7,
                "check_nan_var_names": ["loss"], "loss_names": ["loss"], "scale_sparse_gradient_with_batch_size": False,
                "user_define_dump_filename": "x", "mpi_rank": 1, "mpi_size": 2, "trainers": [0, 1], "trainer_id": 1,
                "adjust_ins_weight": {"need_adjust": True, "nid_slot": "6048"},
                "copy_table": {"need_copy": True, "src_sparse_tables": [0], "dest_sparse_tables": [1], "dependency_map": {"a": "b"}}}
    t = TrainerFactory()._create_trainer(opt_info)
    assert isinstance(t, BoxPSTrainer) and isinstance(t._device_worker, BoxPSWorker)
    t._set_program(model)
    t._set_thread(2)
    t._set_debug(True)
    t._set_fetch_var_and_info(["loss"], ["loss="], 10)
    t._gen_trainer_desc()
    d = parse_text_format(str(t))
    assert d["class_name"] == "BoxPSTrainer" and d["device_worker_name"] == "BoxPSWorker" and d["thread_num"] == 2
    assert d["boxps_param"] == {"sync_dense_mode": 1, "sync_weight_step": 4, "async_mode": False, "dump_thread_num": 3}
    assert d["adjust_ins_weight_config"]["nid_slot"] == "6048" and d["copy_table_config"]["table_denpendency_map"]["key"] == "a"
    assert d["fetch_config"]["print_period"] == 10
    cfg = t.to_config()
    assert isinstance(cfg, TrainerConfig) and cfg.debug and cfg.dump_fields == ["pred", "label"] and cfg.dump_mode == 1
    assert cfg.dump_interval == 7 and cfg.sync_dense_mode == "kstep" and cfg.sync_weight_step == 4 and cfg.dump_thread_num == 3
    assert cfg.check_nan_var_names == ["loss"] and cfg.scale_sparse_gradient_with_batch_size is False
    assert cfg.user_define_dump_filename == "x" and not cfg.is_test

    with pytest.raises(ValueError):
        TrainerFactory()._create_trainer({"trainer": "NoSuchTrainer"})
    with pytest.raises(ValueError):
        DeviceWorkerFactory()._create_device_worker("nosuch")
    with pytest.raises(ValueError):
        MultiTrainer()._set_copy_table_config({"src_var_list": ["a"], "dest_var_list": []})


def test_downpour_worker_and_optimizer_config_on_the_program():
    from paddlebox_b200.optimizer import BoxPSOptimizer, DenseOptimizerSpec
    model = torch.nn.Linear(2, 1)
    BoxPSOptimizer(DenseOptimizerSpec("sgd", 0.05)).minimize(model)
    t = TrainerFactory()._create_trainer({"trainer": "DistMultiTrainer", "device_worker": "DownpourSGD", "fleet_desc": {"x": 1}})
    assert isinstance(t._device_worker, DownpourSGD)
    t._set_program(model)
    t._set_infer(True)
    t._gen_trainer_desc()
    d = parse_text_format(t._desc())
    assert d["device_worker_name"] == "DownpourWorker" and d["downpour_param"]["push_sparse"] is False and d["fleet_desc"] == {"x": 1}
    cfg = t.to_config()
    assert cfg.dense_lr == 0.05 and cfg.dense_optimizer == "sgd" and cfg.is_test


def test_data_feed_desc_file(tmp_path):
    p = tmp_path / "data.proto"
    p.write_text('name: "MultiSlotDataFeed"\nbatch_size: 2\nmulti_slot_desc {\n'
                 '  slots { name: "words" type: "uint64" is_dense: false is_used: false }\n'
                 '  slots { name: "label" type: "uint64" is_dense: false is_used: false }\n}\n')
    d = DataFeedDesc(str(p))
    d.set_batch_size(128)
    d.set_dense_slots(["label"])
    d.set_use_slots(["words", "label"])
    back = parse_text_format(d.desc())
    assert back["batch_size"] == 128 and back["pipe_command"] == "cat"
    s = back["multi_slot_desc"]["slots"]
    assert [x["is_used"] for x in s] == [True, True] and [x["is_dense"] for x in s] == [False, True]
    from paddlebox_b200.dataset import DatasetFactory
    ds = DatasetFactory().create_dataset("InMemoryDataset")
    ds.set_data_feed_desc(d.desc())
    assert ds.batch_size == 128 and [v.name for v in ds.use_vars] == ["words", "label"]
    p.write_text('name: "OtherFeed"\n')
    with pytest.raises(ValueError):
        DataFeedDesc(str(p)).set_use_slots(["a"])


def test_executor_reads_a_program_opt_dict_through_the_factory():
    from paddlebox_b200.box_wrapper import BoxWrapper
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import Executor
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"))
    box.initialize_gpu_and_load_model({})
    model = CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(4,))
    model._pipeline_opt = {"sync_dense_mode": 1, "sync_weight_step": 3, "dump_thread_num": 2}
    model._fleet_opt = {"trainer": "BoxPSTrainer", "device_worker": "BoxPSWorker", "loss_names": ["loss"], "dump_fields_path": "/tmp/x"}
    tr = Executor()._trainer_for(model)
    assert tr.cfg.sync_dense_mode == "kstep" and tr.cfg.sync_weight_step == 3 and tr.cfg.dump_thread_num == 2
    assert tr.cfg.loss_names == ["loss"] and tr.cfg.dump_fields_path == "/tmp/x"
    BoxWrapper.reset_instance()
