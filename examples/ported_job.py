#!/usr/bin/env python
"""A job script in the shape PaddleBox users write them — role maker, strategy object, the UN-fused layer chain
(``_pull_box_sparse`` -> ``sequence_pool`` -> ``continuous_value_model`` -> fc), the ``auc`` layer with
``fleet.metrics.auc`` over the workers, and an epoch loop that resumes from its last checkpoint:

    python -m paddlebox_b200.launch --nproc_per_node 2 --backend gloo examples/ported_job.py --cpu --out /tmp/job
    (run it again: the finished epochs are skipped)

The fused path (``models.ctr`` + ``sparse_embedding_pool``) is what a new job should use; this file shows that a
script written against the reference's layer names keeps its structure.
"""
from __future__ import annotations

import argparse
import os
import sys

import torch
from torch import nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from examples.train_ctr import write_pass_files                        # noqa: E402
from paddlebox_b200 import fleet, layers                                # noqa: E402
from paddlebox_b200.auto_checkpoint import ModelCheckpoint, train_epoch_range   # noqa: E402
from paddlebox_b200.box_wrapper import BoxWrapper                       # noqa: E402
from paddlebox_b200.dataset import DatasetFactory, dense_slot, sparse_slot   # noqa: E402
from paddlebox_b200.optimizer import Adam                               # noqa: E402
from paddlebox_b200.role_maker import PaddleCloudRoleMaker              # noqa: E402
from paddlebox_b200.trainer import CTRTrainer                           # noqa: E402


class Net(nn.Module):
    """slots -> pull -> sum-pool per slot -> CVM (log show, log ctr) -> concat with dense -> fc -> sigmoid."""

    def __init__(self, slots: int, dim: int, dense: int):
        super().__init__()
        self.slots, self.dense = slots, dense
        self.fc = nn.Sequential(nn.Linear(slots * (3 + dim) + dense, 32), nn.ReLU(), nn.Linear(32, 1))
        self.auc = layers.Auc(num_thresholds=4095, slide_steps=20)

    def forward(self, batch):
        label, dense = batch.dense[:, :1], batch.dense[:, 1:1 + self.dense]
        emb = layers._pull_box_sparse(batch)                               # [N, 3 + dim], slot-major
        show_clk = torch.cat([torch.ones_like(label), label], 1)
        cols = []
        for s, e in enumerate(layers.split_by_slot(emb, batch, pad_empty=False)):
            pooled = layers.sequence_pool(e, batch.slot_lod(s), "sum")
            cols.append(layers.continuous_value_model(pooled, show_clk, use_cvm=True))
        logit = self.fc(torch.cat(cols + [dense], 1))
        pred = torch.sigmoid(logit)
        loss = layers.log_loss(pred, label).mean()
        self.auc(pred, label)
        return loss, {"pred": pred.view(-1), "label": label.view(-1)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--slots", type=int, default=4)
    ap.add_argument("--dim", type=int, default=8)
    ap.add_argument("--lines", type=int, default=600)
    ap.add_argument("--cpu", action="store_true")
    a = ap.parse_args()

    strategy = fleet.DistributedStrategy()
    strategy.a_sync = False                                                # dense allreduce every step
    strategy.fleet_desc_configs = {"sparse_learning_rate": 0.1, "sparse_embedx_threshold": 1, "sparse_initial_range": 0.01}
    fleet.init(PaddleCloudRoleMaker(is_collective=True) if "PADDLE_TRAINER_ENDPOINTS" in os.environ else None,
               strategy=strategy)
    rank, world = fleet.worker_index(), fleet.worker_num()
    device = torch.device("cpu") if a.cpu or not torch.cuda.is_available() else torch.device("cuda", torch.cuda.current_device())
    box = BoxWrapper.set_instance(embedx_dim=a.dim, device=device, rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({})
    optimizer = fleet.distributed_optimizer(Adam(learning_rate=5e-3))     # applies the strategy's sparse config too
    torch.manual_seed(0)
    net = Net(a.slots, a.dim, dense=2).to(device)
    trainer = CTRTrainer(net, box, optimizer.minimize(net))

    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var([dense_slot("label", 1), dense_slot("dense", 2)] + [sparse_slot(f"s{i}") for i in range(a.slots)])
    ds.set_batch_size(100)
    ds.set_thread(2)
    data = os.path.join(a.out, "data")
    if rank == 0:
        os.makedirs(data, exist_ok=True)
    fleet.barrier_worker()
    # the sparse table is rank-sharded: every rank checkpoints its own shard (and its replica of the dense part)
    ckpt = [ModelCheckpoint(trainer=trainer, box=box)]
    done = []
    for epoch in train_epoch_range(a.epochs, save_checkpoint_inter=0, path=os.path.join(a.out, "ckpt"), objects=ckpt,
                                   trainer_id=rank):
        if rank == 0:
            write_pass_files(data, n_files=world, lines=a.lines, n_sparse=a.slots, n_dense=2, seed=epoch, key_space=60)
        fleet.barrier_worker()
        files = sorted(os.path.join(data, f) for f in os.listdir(data) if f.startswith(f"part-{epoch:03d}-"))
        ds.set_filelist(fleet.util.get_file_shard(files))
        ds.load_into_memory()
        ds.begin_pass()
        net.auc.reset()
        stats = trainer.train_from_dataset(ds)
        ds.end_pass(False)
        ds.release_memory()
        job_auc = fleet.metrics.auc(net.auc.stat_pos, net.auc.stat_neg)    # over all workers
        n_ins = int(fleet.metrics.sum([stats["instances"]])[0])
        done.append(epoch)
        fleet.util.print_on_rank(f"epoch {epoch}: {n_ins} instances, auc {job_auc:.4f}", 0)
    fleet.util.print_on_rank(f"epochs run: {done}", 0)
    fleet.stop_worker()


if __name__ == "__main__":
    main()
