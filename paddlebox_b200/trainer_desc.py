"""``TrainerDesc`` / ``DeviceWorker`` / ``TrainerFactory``: how the reference's executor describes a training run.

Reference: ``python/paddle/fluid/trainer_desc.py`` (``TrainerDesc`` :25-296 and its subclasses ``MultiTrainer`` /
``DistMultiTrainer`` / ``HeterXpuTrainer`` / ``PSGPUTrainer`` / ``HeterPipelineTrainer`` / ``PipelineTrainer`` /
``BoxPSTrainer`` :298-460), ``device_worker.py`` (``Hogwild`` / ``DownpourLite`` / ``DownpourSGD`` / ``DownpourSGDOPT``
/ ``Section`` / ``HeterSection`` / ``BoxPSWorker``, ``DeviceWorkerFactory`` :24-660), ``trainer_factory.py``
(``TrainerFactory._create_trainer`` :34-125), ``framework/trainer_desc.proto``.  There the Python side fills a
protobuf that the C++ trainer reads; here the same setters fill a plain dict, ``_desc()`` gives the protobuf TEXT of
it (so a desc written by the reference's tooling and one written here are interchangeable through
``TrainerConfig.from_desc``), and ``to_config()`` is the ``TrainerConfig`` the trainer of this framework runs with.
One trainer class runs all of them (one process per GPU, the sparse table in HBM): the ``class_name`` /
``device_worker_name`` are recorded, not dispatched on.
"""
from __future__ import annotations

import os

from .fleet_desc import format_text

_ENUMS = ()


class DeviceWorker:
    """What runs the program on one device; fills the worker part of the trainer description."""
    name = "DeviceWorker"
    param_key = None                                   # the sub-message of TrainerDesc this worker owns

    def __init__(self):
        self._program = None
        self._infer = None
        self._fleet_desc = None

    def _set_infer(self, infer: bool = False) -> None:
        self._infer = bool(infer)

    def _set_fleet_desc(self, fleet_desc) -> None:
        self._fleet_desc = fleet_desc

    def _set_program(self, program) -> None:
        self._program = program

    def _gen_worker_desc(self, trainer_desc: "TrainerDesc") -> None:
        trainer_desc.proto_desc["device_worker_name"] = self.name + ("Worker" if not self.name.endswith("Worker") else "")
        if self.param_key:
            trainer_desc.proto_desc.setdefault(self.param_key, {})


class Hogwild(DeviceWorker):
    """Lock-free workers of one machine (device_worker.py:75-202)."""
    name = "Hogwild"
    param_key = "hogwild_param"

    def _gen_worker_desc(self, trainer_desc):
        super()._gen_worker_desc(trainer_desc)
        if self._infer:
            trainer_desc.proto_desc["hogwild_param"]["skip_ops"] = ["feed", "push_sparse", "push_sparse_v2", "push_dense",
                                                                     "distributed_push_sparse", "send"]


class _Downpour(DeviceWorker):
    param_key = "downpour_param"

    def _gen_worker_desc(self, trainer_desc):
        super()._gen_worker_desc(trainer_desc)
        # the parameter-server tables of the job: kept as given (fleet_desc text / dict), the sparse table's
        # accessor configures the HBM table through PSGPUWrapper.init_gpu_server / fleet.distributed_optimizer
        if self._fleet_desc is not None:
            trainer_desc.proto_desc["fleet_desc"] = self._fleet_desc if isinstance(self._fleet_desc, dict) else \
                str(self._fleet_desc)
        if self._infer:
            trainer_desc.proto_desc["downpour_param"]["push_sparse"] = False
            trainer_desc.proto_desc["downpour_param"]["push_dense"] = False


class DownpourLite(_Downpour):
    name = "DownpourLite"


class DownpourSGD(_Downpour):
    name = "DownpourSGD"

    def _gen_worker_desc(self, trainer_desc):
        super()._gen_worker_desc(trainer_desc)
        trainer_desc.proto_desc["device_worker_name"] = "DownpourWorker"


class DownpourSGDOPT(_Downpour):
    name = "DownpourSGDOPT"

    def _gen_worker_desc(self, trainer_desc):
        super()._gen_worker_desc(trainer_desc)
        trainer_desc.proto_desc["device_worker_name"] = "DownpourWorkerOpt"


class Section(DeviceWorker):
    """A pipeline section (device_worker.py:545-592); ``program._pipeline_opt`` carries the schedule."""
    name = "Section"
    param_key = "section_param"

    def _gen_worker_desc(self, trainer_desc):
        super()._gen_worker_desc(trainer_desc)
        opt = getattr(self._program, "_pipeline_opt", None)
        if isinstance(opt, dict):
            sp = trainer_desc.proto_desc["section_param"]
            for k in ("num_microbatches", "start_cpu_core_id", "pipeline_stage", "num_pipeline_stages", "schedule_mode"):
                if k in opt:
                    sp[k] = opt[k]


class HeterSection(Section):
    name = "HeterSection"
    param_key = "heter_section_param"

    def _gen_worker_desc(self, trainer_desc):
        DeviceWorker._gen_worker_desc(self, trainer_desc)


class BoxPSWorker(DeviceWorker):
    """The PaddleBox worker (device_worker.py:623-653): dense sync mode, async mode, dump threads from
    ``program._pipeline_opt``."""
    name = "BoxPSWorker"
    param_key = "boxps_param"

    def _gen_worker_desc(self, trainer_desc):
        super()._gen_worker_desc(trainer_desc)
        opt = getattr(self._program, "_pipeline_opt", None)
        bp = trainer_desc.proto_desc["boxps_param"]
        if isinstance(opt, dict):
            for k in ("sync_dense_mode", "sync_weight_step", "async_mode", "dump_thread_num", "sync_one_ring", "start_cpu_core_id"):
                if k in opt:
                    bp[k] = opt[k]
        elif opt is not None:                          # a TrainerConfig attached by BoxPSOptimizer.minimize
            bp["async_mode"] = bool(opt.async_mode)
            bp["sync_weight_step"] = int(opt.sync_weight_step)
            bp["dump_thread_num"] = int(opt.dump_thread_num)
            bp["sync_dense_mode"] = {"allreduce": 0, "kstep": 1}.get(opt.sync_dense_mode, 0)


class DeviceWorkerFactory:
    def _create_device_worker(self, worker_type: str) -> DeviceWorker:
        classes = {c.__name__.lower(): c for c in (Hogwild, DownpourLite, DownpourSGD, DownpourSGDOPT, Section,
                                                   HeterSection, BoxPSWorker)}
        try:
            return classes[worker_type.lower()]()
        except KeyError:
            raise ValueError(f"unknown device worker {worker_type!r}") from None


class TrainerDesc:
    """The run description: ``proto_desc`` holds the fields of ``trainer_desc.proto`` as a dict."""
    class_name = "MultiTrainer"

    def __init__(self):
        self.proto_desc: dict = {"thread_num": os.cpu_count() or 1}
        self._fleet_desc = None
        self._device_worker: DeviceWorker | None = None
        self._program = None
        self._infer = False

    # ---- setters (the reference's names) -----------------------------------------------------------------------
    def _set_fetch_var_and_info(self, fetch_vars, fetch_info, print_period) -> None:
        fc = self.proto_desc.setdefault("fetch_config", {"fetch_var_names": [], "fetch_var_str_format": []})
        for v, info in zip(fetch_vars, list(fetch_info)):
            fc["fetch_var_names"].append(getattr(v, "name", v))
            fc["fetch_var_str_format"].append(info)
        fc["print_period"] = int(print_period)

    def _set_debug(self, debug) -> None:
        self.proto_desc["debug"] = bool(debug)

    def _set_thread(self, thread_num) -> None:
        self.proto_desc["thread_num"] = int(thread_num)

    def _set_device_worker(self, device_worker: DeviceWorker) -> None:
        self._device_worker = device_worker

    def _set_infer(self, infer) -> None:
        self._infer = bool(infer)

    def _set_fleet_desc(self, fleet_desc) -> None:
        self._fleet_desc = fleet_desc

    def _set_program(self, program) -> None:
        self._program = program

    def _set_trainer_id(self, trainer_id) -> None:
        self.proto_desc["trainer_id"] = int(trainer_id)

    def _set_trainers(self, trainers) -> None:
        self.proto_desc["trainers"] = [int(t) for t in trainers]

    def _set_use_cvm(self, use_cvm=False) -> None:
        self.proto_desc["use_cvm"] = bool(use_cvm)

    def _set_no_cvm(self, no_cvm=False) -> None:
        self.proto_desc["no_cvm"] = bool(no_cvm)

    def _set_scale_sparse_grad_with_batch_size(self, scale_sparse_gradient_with_batch_size=True) -> None:
        self.proto_desc["scale_sparse_gradient_with_batch_size"] = bool(scale_sparse_gradient_with_batch_size)

    def _set_scale_datanorm(self, scale_datanorm=-1) -> None:
        self.proto_desc["scale_datanorm"] = scale_datanorm

    def _set_dump_slot(self, dump_slot) -> None:
        self.proto_desc["dump_slot"] = bool(dump_slot)

    def _set_mpi_rank(self, mpi_rank) -> None:
        self.proto_desc["mpi_rank"] = int(mpi_rank)

    def _set_mpi_size(self, mpi_size) -> None:
        self.proto_desc["mpi_size"] = int(mpi_size)

    def _set_dump_fields(self, dump_fields) -> None:
        self.proto_desc.setdefault("dump_fields", []).extend(str(f) for f in dump_fields)

    def _set_is_dump_in_simple_mode(self, is_dump_in_simple_mode) -> None:
        self.proto_desc["is_dump_in_simple_mode"] = bool(is_dump_in_simple_mode)

    def _set_dump_fields_path(self, path) -> None:
        self.proto_desc["dump_fields_path"] = str(path)

    def _set_dump_file_num(self, dump_file_num) -> None:
        self.proto_desc["dump_file_num"] = int(dump_file_num)

    def _set_user_define_dump_filename(self, user_define_dump_filename) -> None:
        self.proto_desc["user_define_dump_filename"] = str(user_define_dump_filename)

    def _set_dump_converter(self, converter) -> None:
        self.proto_desc["dump_converter"] = str(converter)

    def _set_enable_random_dump(self, enable_random_dump) -> None:
        self.proto_desc["enable_random_dump"] = bool(enable_random_dump)

    def _set_dump_interval(self, dump_interval) -> None:
        self.proto_desc["dump_interval"] = int(dump_interval)

    def _set_random_with_lineid(self, random_with_lineid) -> None:
        self.proto_desc["random_with_lineid"] = bool(random_with_lineid)

    def _set_dump_param(self, dump_param) -> None:
        self.proto_desc.setdefault("dump_param", []).extend(str(p) for p in dump_param)

    def _set_worker_places(self, worker_places) -> None:
        self.proto_desc.setdefault("worker_places", []).extend(int(p) for p in worker_places)

    def _set_use_ps_gpu(self, use_ps_gpu=False) -> None:
        self.proto_desc["use_ps_gpu"] = bool(use_ps_gpu)

    def _set_thread_barrier(self, thread_barrier) -> None:
        self.proto_desc["thread_barrier"] = bool(thread_barrier)

    def _set_check_nan_var_names(self, check_nan_var_names) -> None:
        self.proto_desc.setdefault("check_nan_var_names", []).extend(str(v) for v in check_nan_var_names)

    def _set_loss_names(self, loss_names) -> None:
        self.proto_desc.setdefault("loss_names", []).extend(str(v) for v in loss_names)

    def _set_adjust_ins_weight(self, config_dict) -> None:
        self.proto_desc["adjust_ins_weight_config"] = {
            "need_adjust": bool(config_dict.get("need_adjust", False)), "nid_slot": config_dict.get("nid_slot", ""),
            "nid_adjw_threshold": float(config_dict.get("nid_adjw_threshold", 0.0)),
            "nid_adjw_ratio": float(config_dict.get("nid_adjw_ratio", 0.0)),
            "ins_weight_slot": config_dict.get("ins_weight_slot", "")}

    def _set_copy_table_config(self, config_dict) -> None:
        def pair(a, b):
            x, y = config_dict.get(a, []), config_dict.get(b, [])
            x, y = (x if isinstance(x, list) else [x]), (y if isinstance(y, list) else [y])
            if len(x) != len(y):
                raise ValueError(f"len({a}) != len({b}), {len(x)} vs {len(y)}")
            return x, y
        cfg = {"need_copy": bool(config_dict.get("need_copy", False)), "batch_num": int(config_dict.get("batch_num", 100))}
        cfg["src_sparse_tables"], cfg["dest_sparse_tables"] = pair("src_sparse_tables", "dest_sparse_tables")
        cfg["src_dense_tables"], cfg["dest_dense_tables"] = pair("src_dense_tables", "dest_dense_tables")
        cfg["src_var_list"], cfg["dest_var_list"] = pair("src_var_list", "dest_var_list")
        deps = []
        for key, values in config_dict.get("dependency_map", {}).items():
            values = values if isinstance(values, list) else [values]
            if len(values) != 1:
                raise ValueError(f"dependency len {len(values)} != 1")
            deps.append({"key": key, "values": list(values)})
        cfg["table_denpendency_map"] = deps
        cfg["dense_pull_after_copy"] = bool(config_dict.get("dense_pull_after_copy", True))
        cfg["enable_dependency"] = bool(config_dict.get("enable_dependency", False))
        cfg["sparse_copy_by_feasign"] = bool(config_dict.get("sparse_copy_by_feasign", True))
        self.proto_desc["copy_table_config"] = cfg

    # ---- results -------------------------------------------------------------------------------------------------
    def _gen_trainer_desc(self) -> None:
        self.proto_desc["class_name"] = self.class_name
        if self._device_worker is not None:
            self._device_worker._set_infer(self._infer)
            self._device_worker._set_program(self._program)
            if self._fleet_desc is not None and self._device_worker._fleet_desc is None:
                self._device_worker._set_fleet_desc(self._fleet_desc)
            self._device_worker._gen_worker_desc(self)

    def _desc(self) -> str:
        """The description as protobuf text (the reference hands the serialized message to C++; the text form is
        what is readable on both sides)."""
        d = dict(self.proto_desc)
        if isinstance(d.get("fleet_desc"), str):       # an embedded prototxt stays out of the trainer text
            d.pop("fleet_desc")
        return format_text(d)

    __str__ = _desc

    def to_config(self):
        """The ``TrainerConfig`` this description means for ``CTRTrainer``."""
        from .trainer import TrainerConfig
        cfg = TrainerConfig.from_desc(self._desc())
        opt = getattr(self._program, "_pipeline_opt", None)
        if opt is not None and not isinstance(opt, dict):         # dense optimizer settings ride on the program
            cfg.dense_lr, cfg.dense_optimizer = opt.dense_lr, opt.dense_optimizer
            if "boxps_param" not in self.proto_desc:
                cfg.sync_dense_mode, cfg.sync_weight_step, cfg.async_mode = opt.sync_dense_mode, opt.sync_weight_step, opt.async_mode
        cfg.is_test = bool(self._infer)
        return cfg


class MultiTrainer(TrainerDesc):
    """Threads of one process (trainer_desc.py:298-318)."""
    class_name = "MultiTrainer"


class DistMultiTrainer(TrainerDesc):
    """Workers against a parameter server (trainer_desc.py:320-342)."""
    class_name = "DistMultiTrainer"


class HeterXpuTrainer(TrainerDesc):
    class_name = "HeterXpuTrainer"


class PSGPUTrainer(TrainerDesc):
    """HeterPS workers (trainer_desc.py:368-390); the runnable counterpart is ``ps_gpu.PSGPUTrainer``."""
    class_name = "PSGPUTrainer"


class HeterPipelineTrainer(TrainerDesc):
    class_name = "HeterPipelineTrainer"


class PipelineTrainer(TrainerDesc):
    class_name = "PipelineTrainer"


class BoxPSTrainer(TrainerDesc):
    """The PaddleBox trainer (trainer_desc.py:439-460)."""
    class_name = "BoxPSTrainer"


class TrainerFactory:
    """``_create_trainer(opt_info)``: the description of a run from a program's ``_fleet_opt`` /
    ``_pipeline_opt`` dict (trainer_factory.py:34-125).  Without ``opt_info``: ``MultiTrainer`` + ``Hogwild``."""

    _TRAINERS = {c.__name__: c for c in (MultiTrainer, DistMultiTrainer, HeterXpuTrainer, PSGPUTrainer,
                                         HeterPipelineTrainer, PipelineTrainer, BoxPSTrainer)}
    _SETTERS = (("trainers", "_set_trainers"), ("trainer_id", "_set_trainer_id"), ("dump_slot", "_set_dump_slot"),
                ("mpi_rank", "_set_mpi_rank"), ("mpi_size", "_set_mpi_size"), ("dump_fields", "_set_dump_fields"),
                ("dump_fields_path", "_set_dump_fields_path"), ("dump_file_num", "_set_dump_file_num"),
                ("user_define_dump_filename", "_set_user_define_dump_filename"), ("dump_converter", "_set_dump_converter"),
                ("dump_param", "_set_dump_param"), ("worker_places", "_set_worker_places"), ("use_ps_gpu", "_set_use_ps_gpu"),
                ("is_dump_in_simple_mode", "_set_is_dump_in_simple_mode"), ("enable_random_dump", "_set_enable_random_dump"),
                ("dump_interval", "_set_dump_interval"), ("random_with_lineid", "_set_random_with_lineid"),
                ("use_cvm", "_set_use_cvm"), ("no_cvm", "_set_no_cvm"),
                ("scale_sparse_gradient_with_batch_size", "_set_scale_sparse_grad_with_batch_size"),
                ("scale_datanorm", "_set_scale_datanorm"), ("adjust_ins_weight", "_set_adjust_ins_weight"),
                ("copy_table", "_set_copy_table_config"), ("check_nan_var_names", "_set_check_nan_var_names"),
                ("loss_names", "_set_loss_names"), ("thread_barrier", "_set_thread_barrier"))

    def _create_trainer(self, opt_info=None) -> TrainerDesc:
        if not opt_info:
            trainer, worker = MultiTrainer(), Hogwild()
        else:
            tname, wname = opt_info.get("trainer", "MultiTrainer"), opt_info.get("device_worker", "Hogwild")
            if tname not in self._TRAINERS:
                raise ValueError(f"unknown trainer {tname!r}")
            trainer, worker = self._TRAINERS[tname](), DeviceWorkerFactory()._create_device_worker(wname)
            for key, setter in self._SETTERS:
                if opt_info.get(key) is not None:
                    getattr(trainer, setter)(opt_info[key])
            if opt_info.get("fleet_desc") is not None:
                worker._set_fleet_desc(opt_info["fleet_desc"])
                trainer._set_fleet_desc(opt_info["fleet_desc"])
        trainer._set_device_worker(worker)
        return trainer
