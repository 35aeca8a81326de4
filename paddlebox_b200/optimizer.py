"""``BoxPSOptimizer`` and the collective "transpiler" of the reference, re-expressed.

Reference: ``fluid.optimizer.BoxPSOptimizer`` (python/paddle/fluid/optimizer.py:7315-7611)
wraps a dense optimizer, optionally splits the program by ``cut_list`` and stamps
``program._pipeline_opt = {trainer: BoxPSTrainer, device_worker: BoxPSWorker,
param_need_sync: [...]}``; the collective transpiler
(python/paddle/fluid/transpiler/collective.py:421-740) rewrites the program with
``c_comm_init_all`` + per-grad / fused ``c_allreduce_sum`` ops.  With PyTorch autograd there
is no program to rewrite: ``minimize`` attaches a ``TrainerConfig`` to the model, and the
data-parallel collectives are chosen by ``sync_dense_mode``.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from .trainer import CTRTrainer, TrainerConfig


@dataclass
class DenseOptimizerSpec:
    kind: str = "adam"
    lr: float = 1e-3
    beta1: float = 0.9
    beta2: float = 0.999
    epsilon: float = 1e-8
    weight_decay: float = 0.0


def Adam(learning_rate=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-8, **_):
    return DenseOptimizerSpec("adam", learning_rate, beta1, beta2, epsilon)


def SGD(learning_rate=1e-2, **_):
    return DenseOptimizerSpec("sgd", learning_rate)


class BoxPSOptimizer:
    """``BoxPSOptimizer(Adam(...), cut_list=[], place_list=[...], concurrency_list=[...],
    queue_size=30, sync_steps=1, start_cpu_core_id=0)``."""

    def __init__(self, optimizer: DenseOptimizerSpec, cut_list=None, place_list=None, concurrency_list=None,
                 queue_size: int = 30, sync_steps: int = 1, start_cpu_core_id: int = 0, sync_dense_mode: str | int =
                 "allreduce", async_mode: bool = False, sync_one_ring: bool = False, dump_thread_num: int = 1):
        self._optimizer = optimizer
        self._cut_list = cut_list or []
        self._place_list = place_list or []
        self._concurrency_list = concurrency_list or []
        self._queue_size = queue_size
        self._sync_steps = sync_steps
        self._start_cpu_core_id = start_cpu_core_id
        # reference enum: DenseKStepNode=1, DenseKStepALL=2, DenseDataNormal=3 (boxps_worker.cc:393-395)
        modes = {1: "kstep", 2: "kstep", 3: "allreduce", 0: "none"}
        self._mode = modes.get(sync_dense_mode, sync_dense_mode)
        self._async = async_mode
        self._dump_thread_num = dump_thread_num

    def minimize(self, model: torch.nn.Module, startup_program=None, parameter_list=None, no_grad_set=None):
        """Attach the trainer configuration to ``model`` (the "program"); returns the config.
        ``Executor.train_from_dataset(model, dataset)`` picks it up."""
        if self._cut_list:
            # pipeline-by-section (_split_program, optimizer.py:7496-7575): the model is cut after the
            # named children; run the sections with model._section_pipeline.forward_backward(...)
            from .parallel.pipeline import SectionPipeline, split_sequential
            sections = split_sequential(model, self._cut_list)
            places = self._place_list or None
            if places is not None and len(places) != len(sections):
                raise ValueError(f"place_list has {len(places)} entries for {len(sections)} sections")
            model._section_pipeline = SectionPipeline(sections, places, self._concurrency_list or None,
                                                      self._queue_size)
        cfg = TrainerConfig(dense_lr=self._optimizer.lr, dense_optimizer=self._optimizer.kind,
                            sync_dense_mode=self._mode, sync_weight_step=max(1, self._sync_steps),
                            async_mode=self._async, dump_thread_num=self._dump_thread_num)
        if getattr(self, "_strategy", None) is not None:          # fleet.DistributedStrategy: dump fields etc.
            td = self._strategy.trainer_desc_configs
            cfg.dump_fields_path = td["dump_fields_path"] or cfg.dump_fields_path
            cfg.dump_fields = list(td["dump_fields"]) or cfg.dump_fields
            cfg.dump_param = list(td["dump_param"]) or cfg.dump_param
        model._pipeline_opt = cfg
        model._param_need_sync = [n for n, p in model.named_parameters() if p.requires_grad]
        return cfg

    def build_trainer(self, model: torch.nn.Module, box=None) -> CTRTrainer:
        cfg = getattr(model, "_pipeline_opt", None) or self.minimize(model)
        return CTRTrainer(model, box, cfg)


class GradAllReduce:
    """Stand-in for ``transpiler.collective.GradAllReduce / SingleProcessMultiThread /
    MultiThread``: there is nothing to transpile; ``transpile`` records the choice and returns
    the sync mode a ``TrainerConfig`` should use."""

    def __init__(self, nrings: int = 1, fuse_all_reduce: bool = True):
        self.nrings, self.fuse = nrings, fuse_all_reduce
        self.mode = "allreduce"

    def transpile(self, startup_program=None, main_program=None, rank: int = 0, endpoints=None,
                  current_endpoint=None, wait_port=None):
        if main_program is not None and hasattr(main_program, "_pipeline_opt") and main_program._pipeline_opt:
            main_program._pipeline_opt.sync_dense_mode = self.mode
        return self.mode


class LocalSGD(GradAllReduce):
    """K-step model averaging (reference ``LocalSGD`` transpiler / ``sync_weight_step``)."""

    def __init__(self, k_steps: int = 1, **kw):
        super().__init__(**kw)
        self.mode = "kstep"
        self.k_steps = k_steps

    @staticmethod
    def snapshot_name(param_name: str) -> str:
        return param_name + "@SNAPSHOT"

    def transpile(self, startup_program=None, main_program=None, **kw):
        mode = super().transpile(startup_program, main_program, **kw)
        if main_program is not None and getattr(main_program, "_pipeline_opt", None):
            main_program._pipeline_opt.sync_weight_step = self.k_steps
        return mode


class SingleProcessMultiThread(GradAllReduce):
    """``collective_mode = "single_process_multi_thread"``: one process driving all local GPUs with per-grad
    allreduce (transpiler/collective.py:421-470).  Here every GPU has its own process; the sync mode is the
    same per-step allreduce."""


class MultiThread(GradAllReduce):
    """The "box" transpiler (transpiler/collective.py:434-740): ``trans_mode`` = ``all_reduce`` (one
    ``c_allreduce_sum`` per gradient), ``fuse_all_reduce`` (gradients coalesced in groups of 128 first) or
    ``all_gather`` (every GPU gathers all gradients and applies them).  All three give every replica the mean
    gradient per step; the flat gradient buffer here is always "fused", so they map to the same sync mode."""

    def __init__(self, nrings: int = 1, trans_mode: str = "all_reduce"):
        if trans_mode not in ("all_reduce", "fuse_all_reduce", "all_gather"):
            raise ValueError(f"unknown trans_mode {trans_mode!r}")
        super().__init__(nrings, fuse_all_reduce=trans_mode != "all_reduce")
        self.trans_mode = trans_mode
        self.fuse_grad_size_in_num = 128


Collective = GradAllReduce      # the transpilers' common base (transpiler/collective.py:36-255)


class DistributeTranspilerConfig:
    """``fluid.DistributeTranspilerConfig``: ``mode`` must be ``"collective"``; ``collective_mode`` picks the
    data-parallel scheme (``grad_allreduce``, ``local_sgd``, ``single_process_multi_thread``)."""

    def __init__(self):
        self.mode = "collective"
        self.collective_mode = "grad_allreduce"
        self.nccl_comm_num = 1
        self.local_sgd_steps = 1


class DistributeTranspiler:
    """``fluid.DistributeTranspiler(config).transpile(trainer_id, program=..., trainers=..., ...)`` as the
    reference's BoxPS test drives it (test_boxps.py:27-72): picks the collective scheme and stamps it on the
    program's trainer configuration."""

    _MODES = {"grad_allreduce": GradAllReduce, "local_sgd": LocalSGD,
              "single_process_multi_thread": SingleProcessMultiThread}

    def __init__(self, config: DistributeTranspilerConfig | None = None):
        self.config = config or DistributeTranspilerConfig()
        self.impl = None

    def transpile(self, trainer_id: int = 0, program=None, pservers: str = "", trainers="127.0.0.1:6174",
                  sync_mode: bool = True, startup_program=None, current_endpoint: str = "127.0.0.1:6174"):
        if self.config.mode != "collective":
            raise ValueError("only config.mode = 'collective' exists here: the parameter server lives in the GPUs")
        cls = self._MODES.get(self.config.collective_mode)
        if cls is None:
            raise ValueError(f"invalid collective_mode {self.config.collective_mode!r}")
        endpoints = trainers.split(",") if isinstance(trainers, str) else list(trainers)
        if cls is GradAllReduce and len(endpoints) == 1 and self.config.collective_mode == "grad_allreduce":
            # the reference refuses a multi-process scheme with a single endpoint (collective.py:82-88)
            raise ValueError("grad_allreduce needs more than one trainer endpoint")
        self.impl = cls(k_steps=self.config.local_sgd_steps) if cls is LocalSGD else cls(self.config.nccl_comm_num)
        return self.impl.transpile(startup_program, program, rank=trainer_id, endpoints=endpoints,
                                   current_endpoint=current_endpoint)
