#!/usr/bin/env python
"""Online (day / pass) training the way production PaddleBox jobs are scripted — around ``FleetUtil``:

    python examples/online_train.py --cpu --out /tmp/job --days 20240101,20240102 --hours "{0..1}"
    python examples/online_train.py --cpu --out /tmp/job --days 20240101,20240102,20240103   # resumes after the last save

Per pass: the data splits of the pass (``get_online_pass_interval``) are loaded, trained with the streamed
pass lifecycle, global metrics are printed, a delta model is saved and listed in ``xbox_patch_done.txt``; at the
end of a day the table is shrunk, the serving base goes to ``<day>/base`` (``xbox_base_done.txt``) and the batch
model with the dense parameters to ``<day>/0`` (``donefile.txt``).  On start the job looks up the last batch
model and the deltas after it and continues from there — the reference's recovery model.
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from examples.train_ctr import write_pass_files                        # noqa: E402
from paddlebox_b200 import fleet                                        # noqa: E402
from paddlebox_b200.box_wrapper import BoxWrapper                       # noqa: E402
from paddlebox_b200.dataset import DatasetFactory, dense_slot, sparse_slot   # noqa: E402
from paddlebox_b200.fleet_util import FleetUtil, expand_braces          # noqa: E402
from paddlebox_b200.models.ctr import build_model                       # noqa: E402
from paddlebox_b200.trainer import CTRTrainer, TrainerConfig            # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True, help="job output directory (models, donefiles)")
    ap.add_argument("--data", default="", help="data root; synthetic splits are written to <data>/<day>/<split>/ if missing")
    ap.add_argument("--days", default="20240101,20240102")
    ap.add_argument("--hours", default="{0..1}")
    ap.add_argument("--split-interval", type=int, default=30, help="minutes of data per split")
    ap.add_argument("--split-per-pass", type=int, default=2)
    ap.add_argument("--slots", type=int, default=6)
    ap.add_argument("--dense", type=int, default=3)
    ap.add_argument("--dim", type=int, default=8)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--lines", type=int, default=512, help="instances per synthetic split")
    ap.add_argument("--cpu", action="store_true")
    a = ap.parse_args()

    fleet.init()
    rank, world = fleet.worker_index(), fleet.worker_num()
    device = torch.device("cpu") if a.cpu or not torch.cuda.is_available() else torch.device("cuda", torch.cuda.current_device())
    box = BoxWrapper.set_instance(embedx_dim=a.dim, device=device, rank=rank, world_size=world)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 1.0, "learning_rate": 0.1, "mf_learning_rate": 0.1})
    box.init_metric("AucCalculator", "join_auc", "label", "pred", metric_phase=1, bucket_size=100000)
    torch.manual_seed(0)
    model = build_model("deepfm", num_slots=a.slots, embedx_dim=a.dim, dense_dim=a.dense, hidden=(64, 32)).to(device)
    trainer = CTRTrainer(model, box, TrainerConfig(dense_lr=5e-3))
    util = FleetUtil(box=box, trainer=trainer)
    data_root = a.data or os.path.join(a.out, "data")

    # ---- resume: the last batch model, then every delta listed after it ------------------------------------
    last_day, last_pass, last_path, base_key = util.get_last_save_model(a.out)
    if last_day > 0:
        util.rank0_print(f"resume from {last_path} (day {last_day}, pass {last_pass})")
        util.load_model(a.out, last_day, last_pass)
        d_day, d_pass, d_path, _ = util.get_last_save_xbox(a.out)
        if d_day > last_day or (d_day == last_day and last_pass != -1 and d_pass > last_pass):
            util.rank0_print(f"(deltas up to {d_day}/{d_pass} exist; a serving side applies them, training restarts "
                             f"from the batch model)")
    else:
        base_key = int(time.time())

    ds = DatasetFactory().create_dataset("PadBoxSlotDataset")
    ds.set_use_var([dense_slot("label", 1), dense_slot("dense", a.dense)] + [sparse_slot(f"slot{i}") for i in range(a.slots)])
    ds.set_batch_size(a.batch)
    ds.set_thread(2)
    passes = util.get_online_pass_interval(a.days, a.hours, a.split_interval, a.split_per_pass, False)
    summary = []
    for day in expand_braces(a.days):
        if int(day) <= last_day and last_pass == -1:
            continue                                            # this day's batch model exists already
        ds.set_date(day)
        for pass_id, splits in enumerate(passes, 1):
            files = []
            for sp in splits:
                d = os.path.join(data_root, day, sp)
                if not os.path.isdir(d) or not os.listdir(d):
                    os.makedirs(d, exist_ok=True)
                    if rank == 0:
                        write_pass_files(d, world, a.lines, a.slots, a.dense, seed=(int(day) * 131 + int(sp)) % 100003)
                fleet.barrier_worker()
                files += sorted(os.path.join(d, f) for f in os.listdir(d))
            ds.set_filelist(fleet.util.get_file_shard(files))
            ds.load_into_memory()
            ds.begin_pass()
            stats = trainer.train_from_dataset(ds)
            ds.end_pass(False)
            ds.release_memory()
            m = util.print_global_metrics("join_auc", print_prefix=f"{day} pass {pass_id}")
            summary.append((day, pass_id, m[0], stats["instances"]))
            util.save_delta_model(a.out, day, pass_id)
            util.write_xbox_donefile(a.out, day, pass_id, base_key, os.path.join(data_root, day))
        # ---- end of day ------------------------------------------------------------------------------------------
        removed = fleet.shrink_sparse_table(decay=0.98, delete_threshold=0.0, delete_after_unseen_days=30)
        util.rank0_print(f"{day}: shrink removed {removed} rows")
        base_key = int(time.time())
        util.save_xbox_base_model(a.out, day)
        util.write_xbox_donefile(a.out, day, -1, base_key, os.path.join(data_root, day))
        util.save_batch_model(a.out, day)
        util.write_model_donefile(a.out, day, -1, base_key)
    if rank == 0:
        print("summary:", " ".join(f"{d}/{p}:auc={auc:.4f}" for d, p, auc, _ in summary))
        print(f"last model: {util.get_last_save_model(a.out)[:3]}")
    fleet.stop_worker()


if __name__ == "__main__":
    main()
