"""HeterPS-flavoured names over the same engine.

The reference ships two front ends for the GPU parameter server: the closed-library
``BoxWrapper`` and its open twin ``PSGPUWrapper`` / ``pull_gpups_sparse`` / ``PSGPUTrainer``
(fleet/ps_gpu_wrapper.h:1-728, operators/pull_gpups_sparse_op.h, ps_gpu_trainer.cc:37-420).
Both drive one sharded HBM table here, so the twin is a naming layer.
"""
from __future__ import annotations

from .box_wrapper import BoxWrapper
from .ops.pull_box_sparse import pull_box_sparse, sparse_embedding_pool
from .trainer import CTRTrainer, Executor, TrainerConfig


class PSGPUWrapper(BoxWrapper):
    """``PSGPUWrapper.get_instance()`` — same object model as :class:`BoxWrapper`."""

    # PreBuildTask + BuildPull + BuildGPUTask == begin/end feed pass + begin pass
    def build_task(self, keys=None) -> None:
        self.begin_feed_pass()
        if keys is not None:
            self.feed_pass(keys)
        self.end_feed_pass()
        self.begin_pass()

    def load_into_memory(self, dataset) -> None:
        dataset.load_into_memory()

    def set_slot_vector(self, slots) -> None:
        self.slot_vector = list(slots)

    def set_slot_dim_vector(self, dims) -> None:
        """Per-slot embedx widths (dynamic-mf mode).  With more than one distinct width, build a
        ``multi_dim.MultiDimBox(self.slot_dim_vector)`` — one table per width."""
        self.slot_dim_vector = list(dims)

    def set_slot_offset_vector(self, offsets) -> None:
        self.slot_offset_vector = list(offsets)

    def set_dataset(self, dataset) -> None:
        self.dataset = dataset

    def init(self, *a, **kw) -> None:
        """``PSGPUWrapper::init``: nothing to pre-create here (tables are built by ``init_gpu_ps``)."""

    def init_gpu_ps(self, gpus=None, conf: dict | None = None, **kw) -> None:
        """``InitializeGPU(dev_ids)``: this process drives the current device only (one process per GPU)."""
        if not self._initialized:
            self.initialize_gpu_and_load_model(conf or {}, **kw)

    # ---- AFS / HDFS client names of ps_gpu_wrapper_py.cc (delegating to BoxFileMgr) -----------------
    def _fm(self):
        fm = getattr(self, "file_mgr", None)
        if fm is None:
            from .side_tables import BoxFileMgr
            fm = self.file_mgr = BoxFileMgr()
        return fm

    def list(self, path: str) -> list:
        return self._fm().list_dir(path)

    def exist(self, path: str) -> bool:
        return self._fm().exists(path)

    def mkdir(self, path: str) -> bool:
        return self._fm().makedir(path)

    def remove(self, path: str) -> bool:
        return self._fm().remove(path)

    def mv(self, src: str, dst: str) -> bool:
        return self._fm().rename(src, dst)

    def touchz(self, path: str) -> bool:
        return self._fm().touch(path)

    def upload(self, local: str, remote: str) -> bool:
        return self._fm().upload(local, remote)

    def download(self, local: str, remote: str) -> bool:
        return self._fm().download(remote, local)

    def cat(self, path: str) -> str:
        fm = self._fm()
        if fm._remote(path):
            return fm._run("-cat", path).stdout
        with open(path) as f:
            return f.read()

    # SetSparseSGD / SetEmbedxSGD (fleet/ps_gpu_wrapper.cu:139-187): the two halves of the sparse
    # optimizer configuration, applied to the host tier and the HBM shard
    def set_sparse_sgd(self, nonclk_coeff=0.1, clk_coeff=1.0, min_bound=-10.0, max_bound=10.0, learning_rate=0.05,
                       initial_g2sum=3.0, initial_range=0.0, beta1_decay_rate=0.9, beta2_decay_rate=0.999,
                       ada_epsilon=1e-8) -> None:
        PSGPUWrapper._update_sparse_config(self, nonclk_coeff=nonclk_coeff, clk_coeff=clk_coeff, min_bound=min_bound,
                                   max_bound=max_bound, learning_rate=learning_rate, initial_g2sum=initial_g2sum,
                                   initial_range=initial_range, beta1_decay_rate=beta1_decay_rate,
                                   beta2_decay_rate=beta2_decay_rate, ada_epsilon=ada_epsilon)

    def set_embedx_sgd(self, mf_create_thresholds=10.0, mf_learning_rate=0.05, mf_initial_g2sum=3.0,
                       mf_initial_range=1e-4, mf_min_bound=-10.0, mf_max_bound=10.0, mf_beta1_decay_rate=0.9,
                       mf_beta2_decay_rate=0.999, mf_ada_epsilon=1e-8) -> None:
        PSGPUWrapper._update_sparse_config(self, mf_create_thresholds=mf_create_thresholds, mf_learning_rate=mf_learning_rate,
                                   mf_initial_g2sum=mf_initial_g2sum, mf_initial_range=mf_initial_range,
                                   mf_min_bound=mf_min_bound, mf_max_bound=mf_max_bound,
                                   mf_beta1_decay_rate=mf_beta1_decay_rate, mf_beta2_decay_rate=mf_beta2_decay_rate,
                                   mf_ada_epsilon=mf_ada_epsilon)

    def _update_sparse_config(self, **kw) -> None:
        from . import _native as N
        cur = self.host_table.config.as_dict() if getattr(self, "host_table", None) is not None else {}
        cur.update(kw)
        cfg = N.SparseOptConfig(**cur)
        if getattr(self, "host_table", None) is not None:
            self.host_table.set_config(cfg)
        if getattr(self, "table", None) is not None:
            self.table.set_config(cfg)

    def init_gpu_server(self, conf: dict | None = None, **kw) -> None:
        """``InitializeGPUServer``: sparse-SGD config from the fleet desc (ps_gpu_trainer.cc:137-254)."""
        self.initialize_gpu_and_load_model(conf or {}, **kw)


def pull_gpups_sparse(batch, size=None, offset: int = 0, box=None):
    """``_pull_gpups_sparse`` (python/paddle/fluid/layers/nn.py): HeterPS twin of ``pull_box_sparse``."""
    return pull_box_sparse(batch, size=size, offset=offset, box=box)


PSGPUTrainer = CTRTrainer
PSGPUWorker = CTRTrainer
__all__ = ["PSGPUWrapper", "pull_gpups_sparse", "PSGPUTrainer", "PSGPUWorker", "sparse_embedding_pool", "Executor",
           "TrainerConfig"]
