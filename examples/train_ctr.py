#!/usr/bin/env python
"""End-to-end example: day-by-day, pass-by-pass sparse-CTR training the PaddleBox way.

    python examples/train_ctr.py [--model deepfm|ctr_dnn|wide_deep] [--passes 3] [--cpu] [--out DIR]
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/train_ctr.py      # one process per GPU

It writes synthetic slot files in the reference text format (``data_generator``), then for every pass:
preload -> feed pass (keys to their owners) -> begin_pass (host tier -> HBM) -> train -> end_pass (+ delta
model) -> metrics, and at the end of the "day": shrink, batch model + serving (xbox) base model.
The flow is the reference's user script; only the model is a ``torch.nn.Module``.
"""
from __future__ import annotations

import argparse
import os
import random
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from paddlebox_b200.box_wrapper import BoxWrapper                      # noqa: E402
from paddlebox_b200.data_generator import format_instance               # noqa: E402
from paddlebox_b200.dataset import DatasetFactory, dense_slot, sparse_slot   # noqa: E402
from paddlebox_b200.models.ctr import build_model                       # noqa: E402
from paddlebox_b200.trainer import CTRTrainer, TrainerConfig            # noqa: E402
from paddlebox_b200.checkpoint import save_persistables                 # noqa: E402
from paddlebox_b200.utils.donefile import ModelManifest                 # noqa: E402


def write_pass_files(directory: str, n_files: int, lines: int, n_sparse: int, n_dense: int, seed: int,
                     key_space: int = 200) -> list:
    """Clicks depend on the first id of slot 0 so that there is something to learn."""
    rng = random.Random(seed)
    ctr = {k: random.Random(k).random() for k in range(1, key_space + 1)}
    files = []
    for f in range(n_files):
        path = os.path.join(directory, f"part-{seed:03d}-{f:05d}")
        with open(path, "w") as fp:
            for _ in range(lines):
                sparse = [[rng.randint(1, key_space) + s * key_space for _ in range(rng.randint(1, 3))]
                          for s in range(n_sparse)]
                label = 1.0 if rng.random() < ctr[sparse[0][0]] else 0.0
                dense = [round(rng.random(), 4) + 0.001 for _ in range(n_dense)]
                fp.write(format_instance([[label], dense] + sparse) + "\n")
        files.append(path)
    return files


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="deepfm", choices=["deepfm", "ctr_dnn", "wide_deep"])
    ap.add_argument("--passes", type=int, default=3)
    ap.add_argument("--lines", type=int, default=2048, help="instances per file")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--slots", type=int, default=8)
    ap.add_argument("--dense", type=int, default=4)
    ap.add_argument("--dim", type=int, default=8)
    ap.add_argument("--key-space", type=int, default=200, help="distinct ids per slot")
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--out", default="")
    a = ap.parse_args()

    use_cuda = torch.cuda.is_available() and not a.cpu
    rank, world = 0, 1
    if "RANK" in os.environ:                                    # launched by torchrun: one process per GPU
        import torch.distributed as dist
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        if use_cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        dist.init_process_group("nccl" if use_cuda else "gloo")
    device = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")

    box = BoxWrapper.set_instance(embedx_dim=a.dim, device=device, rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 1.0, "learning_rate": 0.1, "mf_learning_rate": 0.1},
                                      table_capacity=1 << 20, max_keys_per_batch=1 << 16)
    box.init_metric("AucCalculator", "auc", "label", "pred", metric_phase=1, bucket_size=100000)

    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var([dense_slot("label", 1), dense_slot("dense", a.dense)] + [sparse_slot(f"s{i}") for i in range(a.slots)])
    ds.set_batch_size(a.batch)
    ds.set_thread(2)
    ds.set_rank_world(rank, world)

    torch.manual_seed(0)
    model = build_model(a.model, num_slots=a.slots, embedx_dim=a.dim, dense_dim=a.dense, hidden=(64, 32)).to(device)
    trainer = CTRTrainer(model, box, TrainerConfig(dense_lr=5e-3))

    out = a.out or tempfile.mkdtemp(prefix="pbx_example_")
    data_dir = os.path.join(out, "data")
    os.makedirs(data_dir, exist_ok=True)
    ds.set_date("20240101")
    manifest = ModelManifest(os.path.join(out, "donefiles"))        # what serving and a restarted job read
    aucs = []
    for p in range(a.passes):
        if rank == 0:                     # one writer; the others wait (a second writer would truncate files being read)
            write_pass_files(data_dir, n_files=2 * world, lines=a.lines, n_sparse=a.slots, n_dense=a.dense, seed=p,
                             key_space=a.key_space)
        if world > 1:
            torch.distributed.barrier()
        files = [os.path.join(data_dir, f"part-{p:03d}-{f:05d}") for f in range(2 * world)]
        ds.set_filelist(files)
        ds.preload_into_memory()          # read + parse + collect keys while (in a real job) the previous pass trains
        ds.wait_preload_done()            # ... and route the keys to their owning ranks (feed pass)
        ds.begin_pass()                   # host tier -> HBM
        stats = trainer.train_from_dataset(ds)
        ds.end_pass(False)                # HBM -> host tier
        box.save_delta(os.path.join(out, f"xbox_delta_{p}"))
        if rank == 0:
            save_persistables(model, os.path.join(out, f"dense_{p}"))
            manifest.record_delta("20240101", p, os.path.join(out, f"xbox_delta_{p}"), os.path.join(out, f"dense_{p}"))
        msg = box.get_metric_msg("auc")
        aucs.append(msg[0])
        if rank == 0:
            print(f"pass {p}: {stats['instances']} instances, {stats['ins_per_sec']:.0f} ins/s, auc {msg[0]:.4f} "
                  f"bucket_error {msg[1]:.4f} mae {msg[2]:.4f} copc {msg[6]:.3f}")
        ds.release_memory()
    kept = box.shrink_table(decay=0.98, delete_threshold=0.0, delete_after_unseen_days=30)
    info = box.save_base(os.path.join(out, "batch_model"), os.path.join(out, "xbox_base"), "20240101")
    if rank == 0:
        save_persistables(model, os.path.join(out, "dense_base"))
        manifest.record_base("20240101", os.path.join(out, "xbox_base"), os.path.join(out, "batch_model"),
                             os.path.join(out, "dense_base"), pass_id=a.passes)
        print(f"day end: {kept} rows kept, {info}; models under {out}")
        print(box.print_sync_timer())
    return aucs


if __name__ == "__main__":
    main()
