"""``fleet.DistributedStrategy``: how a job wants to be distributed, as one object of switches and config dicts.

Reference: ``python/paddle/distributed/fleet/base/distributed_strategy.py`` over
``framework/distributed_strategy.proto``.  The parameter-server half is what PaddleBox / GPUPS jobs set and what has
a meaning here: ``a_sync`` / ``a_sync_configs`` (dense update mode: synchronous, k-step, asynchronous),
``trainer_desc_configs`` (dump fields), ``fleet_desc_configs`` / ``sparse_table_configs`` (the sparse table's
accessor and optimizer), ``adam_d2sum``, ``fs_client_param``.  The collective half (amp, recompute, sharding,
pipeline, tensor parallel ...) configures stock-Paddle program rewrites PaddleBox does not use; those switches are
kept as plain attributes with the reference's defaults so that a job script can set and read them, and they are
listed by ``unused_switches()`` when turned on.  ``fleet.distributed_optimizer(optimizer, strategy)`` applies the
object (``apply_to``).
"""
from __future__ import annotations

import copy

_A_SYNC_KEYS = {"k_steps": -1, "max_merge_var_num": 1, "send_queue_size": 16, "independent_recv_thread": False,
                "min_send_grad_num_before_recv": 1, "thread_pool_size": 1, "send_wait_times": 1,
                "runtime_split_send_recv": False, "launch_barrier": True, "heter_worker_device_guard": "cpu",
                "lr_decay_steps": 10, "use_ps_gpu": 0, "use_gpu_graph": 0}
_TRAINER_DESC_KEYS = {"dump_fields_path": "", "dump_fields": [], "dump_param": [], "stat_var_names": [], "trainer": "",
                      "device_worker": "", "local_sparse": [], "remote_sparse": []}
_FS_KEYS = {"uri": "", "user": "", "passwd": "", "hadoop_bin": ""}
# collective / program-rewrite switches of the reference with their defaults
_SWITCHES = {"amp": False, "recompute": False, "sharding": False, "pipeline": False, "tensor_parallel": False, "localsgd": False,
             "adaptive_localsgd": False, "dgc": False, "lars": False, "lamb": False, "gradient_merge": False,
             "fuse_all_reduce_ops": True, "fuse_grad_size_in_MB": 32, "fuse_grad_merge": False, "nccl_comm_num": 1,
             "sync_nccl_allreduce": True, "use_hierarchical_allreduce": False, "hierarchical_allreduce_inter_nranks": 1,
             "sync_batch_norm": False, "elastic": False, "auto": False, "semi_auto": False, "auto_search": False,
             "cudnn_exhaustive_search": False, "conv_workspace_size_limit": 512, "cudnn_batchnorm_spatial_persistent": False,
             "find_unused_parameters": False, "without_graph_optimization": False, "heter_ccl_mode": False,
             "is_fl_ps_mode": False, "is_with_coordinator": False, "asp": False, "qat": False, "last_comm_group_size_MB": 1,
             "fuse_grad_size_in_num": 8, "calc_comm_same_stream": False}
_CONFIG_DICTS = ("amp_configs", "recompute_configs", "sharding_configs", "pipeline_configs", "tensor_parallel_configs",
                 "localsgd_configs", "adaptive_localsgd_configs", "dgc_configs", "lars_configs", "lamb_configs",
                 "gradient_merge_configs", "hybrid_configs", "gradient_scale_configs", "qat_configs")


def _checked(known: dict, configs: dict, what: str) -> dict:
    bad = [k for k in configs if k not in known]
    if bad:
        raise ValueError(f"key `{bad[0]}` not in {what}; allowed: {sorted(known)}")
    return dict(configs)


class DistributedStrategy:
    def __init__(self):
        object.__setattr__(self, "_v", {
            "a_sync": True, "a_sync_configs": dict(_A_SYNC_KEYS), "trainer_desc_configs": copy.deepcopy(_TRAINER_DESC_KEYS),
            "adam_d2sum": False, "fs_client_param": dict(_FS_KEYS), "sparse_table_configs": {}, "fleet_desc_configs": {},
            **copy.deepcopy(_SWITCHES), **{k: {} for k in _CONFIG_DICTS}})

    def __getattr__(self, name):
        v = object.__getattribute__(self, "_v")
        if name in v:
            return v[name]
        raise AttributeError(f"DistributedStrategy has no field {name!r}")

    def __setattr__(self, name, value):
        v = self._v
        if name not in v:
            raise AttributeError(f"DistributedStrategy has no field {name!r}")
        if name == "a_sync_configs":
            value = {**v[name], **_checked(_A_SYNC_KEYS, value, "a_sync_configs")}
        elif name == "trainer_desc_configs":
            value = {**v[name], **_checked(_TRAINER_DESC_KEYS, value, "trainer_desc_configs")}
        elif name == "fs_client_param":
            value = {**v[name], **_checked(_FS_KEYS, value, "fs_client_param")}
        elif name == "fleet_desc_configs":
            value = self._checked_fleet_desc(value)
        elif name in _CONFIG_DICTS or name == "sparse_table_configs":
            if not isinstance(value, dict):
                raise TypeError(f"{name} must be a dict")
            value = {**v[name], **value}
        elif isinstance(v[name], bool):
            if not isinstance(value, bool):
                raise ValueError(f"{name} should have value of bool type")
        v[name] = value

    @staticmethod
    def _checked_fleet_desc(configs: dict) -> dict:
        """Keys are the pslib strategy keys (``sparse_learning_rate`` ..., optionally prefixed ``embed_`` /
        ``embedx_``), either flat for the one table or per table name (distributed_strategy.py:520-640)."""
        from .fleet_desc import STRATEGY_KEYS
        flat = {}
        for k, val in configs.items():
            items = val.items() if isinstance(val, dict) else [(k, val)]
            for kk, vv in items:
                base = kk
                for pre in ("embedx_", "embed_"):
                    if base.startswith(pre):
                        base = base[len(pre):]
                if base not in STRATEGY_KEYS:
                    raise ValueError(f"strategy key '{kk}' not support")
                flat[kk] = vv
        return flat

    # ---- what it means here ----------------------------------------------------------------------------------------
    def dense_mode(self):
        """(sync_dense_mode, sync_weight_step, async_mode) of the dense part: ``a_sync`` off = allreduce every step;
        on with ``k_steps`` > 0 = local steps with a parameter sync every k (GEO); on with ``k_steps`` <= 0 = the
        asynchronous host-side optimizer."""
        if not self.a_sync:
            return "allreduce", 1, False
        k = int(self.a_sync_configs.get("k_steps", -1))
        return ("kstep", k, False) if k > 0 else ("async", 1, True)

    def unused_switches(self) -> list:
        return [k for k, d in _SWITCHES.items() if isinstance(d, bool) and self._v[k] and not d]

    def apply_to(self, trainer_config) -> None:
        """Fill a ``TrainerConfig``: dense mode and the dump fields."""
        trainer_config.sync_dense_mode, trainer_config.sync_weight_step, trainer_config.async_mode = self.dense_mode()
        td = self.trainer_desc_configs
        if td["dump_fields_path"]:
            trainer_config.dump_fields_path = td["dump_fields_path"]
        if td["dump_fields"]:
            trainer_config.dump_fields = list(td["dump_fields"])
        if td["dump_param"]:
            trainer_config.dump_param = list(td["dump_param"])

    def sparse_strategy(self) -> dict:
        """The pslib strategy dict of the sparse table (``fleet_desc_configs``; ``adam_d2sum`` picks the optimizer)."""
        d = dict(self.fleet_desc_configs)
        if self.adam_d2sum and "sparse_optimizer" not in d:
            d["sparse_optimizer"] = "adam"
        return d

    def save_to_prototxt(self, output: str) -> None:
        from .fleet_desc import format_text
        with open(output, "w") as f:
            f.write(format_text({k: v for k, v in self._v.items() if v not in ({}, [], "")}))

    def load_from_prototxt(self, pb_file: str) -> None:
        from .fleet_desc import parse_text_format
        with open(pb_file) as f:
            for k, val in parse_text_format(f.read()).items():
                if k in self._v:
                    if isinstance(self._v[k], dict) and isinstance(val, dict):
                        self._v[k].update(val)
                    else:
                        self._v[k] = val

    def __repr__(self):
        on = {k: v for k, v in self._v.items() if k not in _SWITCHES or v != _SWITCHES[k]}
        return f"DistributedStrategy({on})"
