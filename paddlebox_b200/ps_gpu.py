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
        self.slot_dim_vector = list(dims)

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
