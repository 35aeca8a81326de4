"""fleet.DistributedStrategy: the switches and config dicts of a job, and what they configure here."""
import pytest
import torch

from paddlebox_b200 import fleet
from paddlebox_b200.box_wrapper import BoxWrapper
from paddlebox_b200.optimizer import Adam


def test_fields_defaults_and_checks(tmp_path):
    s = fleet.DistributedStrategy()
    assert s.a_sync is True and s.a_sync_configs["k_steps"] == -1 and s.amp is False and s.fuse_grad_size_in_MB == 32
    s.a_sync_configs = {"k_steps": 100, "send_queue_size": 32}
    assert s.a_sync_configs["k_steps"] == 100 and s.a_sync_configs["max_merge_var_num"] == 1
    with pytest.raises(ValueError):
        s.a_sync_configs = {"no_such_key": 1}
    with pytest.raises(ValueError):
        s.amp = 1
    with pytest.raises(AttributeError):
        s.no_such_field = 1
    with pytest.raises(AttributeError):
        s.no_such_field
    with pytest.raises(ValueError):
        s.fleet_desc_configs = {"sparse_bogus": 1}
    s.amp = True
    s.amp_configs = {"init_loss_scaling": 1024}
    assert s.unused_switches() == ["amp"] and s.amp_configs == {"init_loss_scaling": 1024}
    assert s.dense_mode() == ("kstep", 100, False)
    s.a_sync_configs = {"k_steps": 0}
    assert s.dense_mode() == ("async", 1, True)
    s.a_sync = False
    assert s.dense_mode() == ("allreduce", 1, False)
    s.trainer_desc_configs = {"dump_fields_path": "/tmp/dump", "dump_fields": ["pred"]}
    s.fs_client_param = {"uri": "afs://x", "user": "u"}
    p = tmp_path / "s.prototxt"
    s.save_to_prototxt(str(p))
    t = fleet.DistributedStrategy()
    t.load_from_prototxt(str(p))
    assert t.a_sync is False and t.amp is True and t.trainer_desc_configs["dump_fields_path"] == "/tmp/dump"
    assert t.fs_client_param["uri"] == "afs://x" and t.a_sync_configs["k_steps"] == 0
    assert "amp" in repr(t)


def test_distributed_optimizer_applies_a_strategy_object():
    box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"))
    box.initialize_gpu_and_load_model({})
    s = fleet.DistributedStrategy()
    s.a_sync_configs = {"k_steps": 8}
    s.trainer_desc_configs = {"dump_fields_path": "/tmp/d", "dump_fields": ["pred", "label"], "dump_param": ["w"]}
    s.fleet_desc_configs = {"sparse_learning_rate": 0.2, "sparse_embedx_threshold": 3, "sparse_delete_threshold": 0.4,
                            "embedding": {"sparse_click_coeff": 2.0}}
    assert s.fleet_desc_configs["sparse_click_coeff"] == 2.0                    # per-table dicts are flattened
    opt = fleet.distributed_optimizer(Adam(0.01), s)
    assert abs(box.host_table.config.learning_rate - 0.2) < 1e-7 and box.host_table.config.mf_create_thresholds == 3
    assert fleet.table_policy["delete_threshold"] == 0.4
    model = torch.nn.Linear(2, 1)
    cfg = opt.minimize(model)
    assert (cfg.sync_dense_mode, cfg.sync_weight_step, cfg.async_mode) == ("kstep", 8, False) and cfg.dense_lr == 0.01
    assert cfg.dump_fields == ["pred", "label"] and cfg.dump_fields_path == "/tmp/d" and cfg.dump_param == ["w"]
    s2 = fleet.DistributedStrategy()
    s2.a_sync = False
    assert fleet.distributed_optimizer(Adam(0.01), s2).minimize(model).sync_dense_mode == "allreduce"
    s3 = fleet.DistributedStrategy()
    cfg3 = fleet.distributed_optimizer(Adam(0.01), s3).minimize(model)
    assert cfg3.async_mode and s3.adam_d2sum is False
    s3.adam_d2sum = True
    assert s3.sparse_strategy()["sparse_optimizer"] == "adam"
    fleet.table_policy.clear()
    BoxWrapper.reset_instance()


def test_strategy_given_to_fleet_init_is_used_by_the_optimizer():
    s = fleet.DistributedStrategy()
    s.a_sync = False
    fleet.init(is_collective=True, strategy=s)
    try:
        opt = fleet.distributed_optimizer(Adam(0.02))
        assert opt._strategy is s and opt.minimize(torch.nn.Linear(2, 1)).sync_dense_mode == "allreduce"
    finally:
        fleet._init_strategy = None
