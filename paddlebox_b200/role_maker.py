"""Role makers: who this process is in the job, read the way the reference's launchers say it.

Reference: ``python/paddle/fluid/incubate/fleet/base/role_maker.py`` (``RoleMakerBase`` :69-188,
``MPISymetricRoleMaker`` :265-478, ``PaddleCloudRoleMaker`` :480-579, ``GeneralRoleMaker`` :581-1005,
``UserDefinedRoleMaker`` :1178-1252, ``UserDefinedCollectiveRoleMaker`` :1254-1306).  A job script written for
PaddleBox does ``fleet.init(GeneralRoleMaker(...))`` under a launcher that exports ``PADDLE_TRAINER_ID`` /
``PADDLE_TRAINER_ENDPOINTS`` / ``TRAINING_ROLE`` (/ ``PADDLE_PSERVERS_IP_PORT_LIST``).  Here every trainer is one
``torch.distributed`` rank: the role maker yields (rank, world size, rendezvous address) — the first trainer
endpoint is the rendezvous — and ``fleet.init(role_maker)`` joins that world (NCCL on GPUs, gloo on CPU).  There
are no parameter-server processes (the table lives in the trainers' HBM): a process launched with
``TRAINING_ROLE=PSERVER`` learns that it is a server (``is_server()``) and has nothing to run; its job script's
``fleet.run_server()`` returns at once.  The collective helpers (``all_reduce_worker`` / ``all_gather`` /
``barrier_worker``) run over the joined world; the reference's HDFS / HTTP key-value rendezvous of
``GeneralRoleMaker`` is replaced by the TCP store of ``torch.distributed`` (same arguments accepted).
"""
from __future__ import annotations

import os

import numpy as np


class Role:
    WORKER = 1
    SERVER = 2
    XPU = 3


class RoleMakerBase:
    def __init__(self):
        self._worker_endpoints: list = []
        self._server_endpoints: list = []
        self._role_is_generated = False
        self._role = None
        self._current_id = -1

    # ---- identity ---------------------------------------------------------------------------------------------
    def generate_role(self) -> None:
        raise NotImplementedError("Please implement this method in child class")

    def _ensure(self) -> None:
        if not self._role_is_generated:
            self.generate_role()

    def is_worker(self) -> bool:
        self._ensure()
        return self._role == Role.WORKER

    def is_server(self) -> bool:
        self._ensure()
        return self._role == Role.SERVER

    def is_first_worker(self) -> bool:
        return self.is_worker() and self.worker_index() == 0

    def worker_num(self) -> int:
        self._ensure()
        return max(1, len(self._worker_endpoints))

    def server_num(self) -> int:
        self._ensure()
        return len(self._server_endpoints)

    def role_id(self) -> int:
        self._ensure()
        return self._current_id

    def worker_index(self) -> int:
        self._ensure()
        return self._current_id if self._role == Role.WORKER else 0

    def server_index(self) -> int:
        self._ensure()
        return self._current_id if self._role == Role.SERVER else 0

    def get_trainer_endpoints(self) -> list:
        self._ensure()
        return list(self._worker_endpoints)

    def get_pserver_endpoints(self) -> list:
        self._ensure()
        return list(self._server_endpoints)

    def to_string(self) -> str:
        self._ensure()
        return f"role: {self._role}, current_id: {self._current_id}, worker_endpoints: {self._worker_endpoints}, " \
               f"server_endpoints: {self._server_endpoints}"

    def rendezvous(self):
        """(master address, master port) = the first trainer endpoint; (None, None) without endpoints."""
        self._ensure()
        if not self._worker_endpoints or ":" not in self._worker_endpoints[0]:
            return None, None
        host, port = self._worker_endpoints[0].rsplit(":", 1)
        return host, int(port)

    # ---- collectives over the joined world ---------------------------------------------------------------------
    @staticmethod
    def _cluster():
        from .parallel.cluster import Cluster
        return Cluster.ins()

    def all_gather(self, input) -> list:
        return self._cluster().allgather_object(input)

    _all_gather = all_gather

    def all_reduce_worker(self, input, output, mode: str = "sum") -> None:
        """``output[:] = reduce(input)`` over the workers; ``mode``: sum / max / min."""
        from .parallel.cluster import OP_MAX, OP_MIN
        a = np.asarray(input, dtype=np.float64)
        c = self._cluster()
        if mode == "sum":
            res = c.allreduce_sum(a.reshape(-1)).reshape(a.shape)
        else:
            op = {"max": OP_MAX, "min": OP_MIN}[mode]
            res = np.array([c.allreduce(float(v), op) for v in a.reshape(-1)]).reshape(a.shape)
        np.asarray(output)[...] = res

    _all_reduce = all_reduce_worker

    def barrier_worker(self) -> None:
        self._cluster().barrier()

    _barrier_worker = barrier_worker
    barrier_all = barrier_worker
    _barrier_all = barrier_worker

    def _finalize(self) -> None:
        pass


def _split(v: str) -> list:
    return [e for e in (v or "").split(",") if e]


class PaddleCloudRoleMaker(RoleMakerBase):
    """The environment the PaddleCloud / ``paddle.distributed.launch`` launchers export."""

    def __init__(self, is_collective: bool = False):
        super().__init__()
        self._is_collective = is_collective

    def generate_role(self) -> None:
        if self._role_is_generated:
            return
        env = os.environ
        if self._is_collective:
            self._current_id = int(env.get("PADDLE_TRAINER_ID", "0"))
            assert env.get("PADDLE_TRAINING_ROLE", "TRAINER") == "TRAINER"
            eps = env.get("PADDLE_TRAINER_ENDPOINTS")
            assert eps is not None, "can't find PADDLE_TRAINER_ENDPOINTS"
            self._worker_endpoints = _split(eps)
            self._role = Role.WORKER
        else:
            try:
                self._server_endpoints = _split(env["PADDLE_PSERVERS_IP_PORT_LIST"])
                n = int(env["PADDLE_TRAINERS_NUM"])
                role = env["TRAINING_ROLE"]
                if role == "TRAINER":
                    self._role, self._current_id = Role.WORKER, int(env["PADDLE_TRAINER_ID"])
                elif role == "PSERVER":
                    self._role = Role.SERVER
                    self._current_id = self._server_endpoints.index(f"{env['POD_IP']}:{env['PADDLE_PORT']}")
                else:
                    raise ValueError("TRAINING_ROLE must be PSERVER or TRAINER")
            except (KeyError, ValueError) as e:
                raise ValueError(f"something wrong with PaddleCloud, please check environment ({e})") from None
            self._worker_endpoints = _split(env.get("PADDLE_TRAINER_ENDPOINTS", "")) or [""] * n
        self._role_is_generated = True


class GeneralRoleMaker(RoleMakerBase):
    """``GeneralRoleMaker(path=..., hdfs_name=..., hdfs_ugi=..., http_ip_port=..., init_timeout_seconds=...)``: the
    role maker PaddleBox / GPUPS jobs use.  Roles come from ``TRAINING_ROLE`` / ``PADDLE_TRAINER_ID`` /
    ``PADDLE_TRAINER_ENDPOINTS`` / ``PADDLE_PSERVERS_IP_PORT_LIST``; the rendezvous arguments are kept (the
    time-out is used for the ``torch.distributed`` rendezvous)."""

    def __init__(self, **kwargs):
        super().__init__()
        self._hdfs_name = kwargs.get("hdfs_name", "")
        self._hdfs_ugi = kwargs.get("hdfs_ugi", "")
        self._hdfs_path = kwargs.get("path", "").rstrip("/")
        self._init_timeout_seconds = kwargs.get("init_timeout_seconds", 3600)
        self._run_timeout_seconds = kwargs.get("run_timeout_seconds", 9999999)
        self._use_metric = kwargs.get("use_metric", False)
        self._use_ps_gpu = kwargs.get("use_ps_gpu", False)
        self._http_ip_port = _split(kwargs.get("http_ip_port", "").replace(":", ","))

    def generate_role(self) -> None:
        if self._role_is_generated:
            return
        env = os.environ
        self._server_endpoints = _split(env.get("PADDLE_PSERVERS_IP_PORT_LIST", ""))
        self._worker_endpoints = _split(env["PADDLE_TRAINER_ENDPOINTS"])
        role = env.get("TRAINING_ROLE", "TRAINER")
        if role == "TRAINER":
            self._role, self._current_id = Role.WORKER, int(env["PADDLE_TRAINER_ID"])
        elif role == "PSERVER":
            self._role = Role.SERVER
            self._current_id = self._server_endpoints.index(f"{env['POD_IP']}:{env['PADDLE_PORT']}")
        else:
            raise ValueError("TRAINING_ROLE must be PSERVER or TRAINER")
        self._role_is_generated = True

    def get_local_endpoint(self) -> str:
        self._ensure()
        return self._worker_endpoints[self._current_id] if self._role == Role.WORKER else \
            self._server_endpoints[self._current_id]


class MPISymetricRoleMaker(RoleMakerBase):
    """Under ``mpirun``: ranks alternate server / worker in the reference (even = server, odd = worker,
    role_maker.py:461-478).  Without server processes every MPI rank is a worker here; rank / size come from
    the MPI launcher's environment (OMPI / PMI) and the rendezvous from ``MASTER_ADDR`` / ``MASTER_PORT``."""

    def generate_role(self) -> None:
        if self._role_is_generated:
            return
        env = os.environ
        rank = int(env.get("OMPI_COMM_WORLD_RANK", env.get("PMI_RANK", env.get("RANK", "0"))))
        size = int(env.get("OMPI_COMM_WORLD_SIZE", env.get("PMI_SIZE", env.get("WORLD_SIZE", "1"))))
        host, port = env.get("MASTER_ADDR", "127.0.0.1"), env.get("MASTER_PORT", "29500")
        self._worker_endpoints = [f"{host}:{port}"] + [f"{host}:0"] * (size - 1)
        self._role, self._current_id = Role.WORKER, rank
        self._role_is_generated = True


class UserDefinedRoleMaker(RoleMakerBase):
    def __init__(self, current_id: int = 0, role: int = Role.WORKER, worker_num: int = 0, server_endpoints=None,
                 worker_endpoints=None):
        super().__init__()
        if role not in (Role.WORKER, Role.SERVER):
            raise TypeError("role must be Role.WORKER or Role.SERVER")
        if current_id < 0:
            raise ValueError("current_id must be greater than or equal to 0")
        self._current_id, self._role = int(current_id), role
        self._server_endpoints = list(server_endpoints or [])
        self._worker_endpoints = list(worker_endpoints or [""] * int(worker_num))
        if role == Role.SERVER and current_id >= max(1, len(self._server_endpoints)):
            raise ValueError("if role is Role.SERVER, current_id must be less than the number of server endpoints")

    def generate_role(self) -> None:
        self._role_is_generated = True


class UserDefinedCollectiveRoleMaker(RoleMakerBase):
    def __init__(self, current_id: int = 0, worker_endpoints=None):
        super().__init__()
        if not worker_endpoints:
            raise ValueError("worker_endpoints must be a non-empty list")
        if not 0 <= current_id < len(worker_endpoints):
            raise ValueError("current_id must be in [0, number of worker endpoints)")
        self._current_id, self._role = int(current_id), Role.WORKER
        self._worker_endpoints = list(worker_endpoints)

    def generate_role(self) -> None:
        self._role_is_generated = True
