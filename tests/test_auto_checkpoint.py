"""Numbered checkpoints (CheckpointSaver) and the resuming epoch loop (train_epoch_range)."""
import os

import numpy as np
import pytest
import torch

from paddlebox_b200.auto_checkpoint import CheckpointSaver, ModelCheckpoint, TrainStatus, train_epoch_range
from paddlebox_b200.box_wrapper import BoxWrapper
from paddlebox_b200.utils.fs import LocalFS


def test_numbered_checkpoints_rename_into_place_and_clean_up(tmp_path):
    root = str(tmp_path / "ckpt")
    saver = CheckpointSaver(LocalFS())
    assert saver.load_checkpoint(root, [TrainStatus()]) is None
    with pytest.raises(FileNotFoundError):
        saver.load_checkpoint(root, [TrainStatus()], ignore_empty=False)
    for e in range(3):
        real, no = saver.save_checkpoint(root, [TrainStatus(e, note="x")])
        assert no == e and real.endswith(f"__paddle_checkpoint__.{e}") and os.path.isdir(real)
    assert saver.get_checkpoint_no(root) == [0, 1, 2] and not [d for d in os.listdir(root) if d.endswith(".tmp")]
    st = TrainStatus()
    assert saver.load_checkpoint(root, [st]).endswith(".2") and st == TrainStatus(2, note="x") and st.next() == 3
    saver.load_checkpoint(root, [st], checkpoint_no=0)
    assert st.epoch_no == 0
    os.makedirs(os.path.join(root, "__paddle_checkpoint__.junk"))            # not a number: ignored
    os.makedirs(os.path.join(root, "other.7"))
    saver.clean_redundant_checkpoints(root)
    assert saver.get_checkpoint_no(root) == [2]
    saver.save_checkpoint(root, [TrainStatus(3)])
    saver.clean_redundant_checkpoints(root, reserved=[2, 3])
    assert saver.get_checkpoint_no(root) == [2, 3]
    (tmp_path / "file").write_text("x")
    with pytest.raises(ValueError):
        saver.save_checkpoint(str(tmp_path / "file"), [TrainStatus()])


class _RemoteFS(LocalFS):
    """A LocalFS that claims to be remote: checkpoints go through the local cache directory."""
    def need_upload_download(self):
        return True


def test_remote_file_system_goes_through_the_cache(tmp_path):
    root, cache = str(tmp_path / "remote"), str(tmp_path / "cache")
    saver = CheckpointSaver(_RemoteFS())
    real, no = saver.save_checkpoint(root, [TrainStatus(5)], trainer_id=1, local_cache_path=cache)
    assert no == 0 and os.path.exists(os.path.join(real, "train_status.json")) and os.listdir(cache) == []
    st = TrainStatus()
    saver.load_checkpoint(root, [st], trainer_id=1, local_cache_path=cache)
    assert st.epoch_no == 5 and os.listdir(cache) == []


def test_epoch_loop_resumes_model_optimizer_and_sparse_table(tmp_path):
    from paddlebox_b200.models.ctr import CtrDnn
    from paddlebox_b200.trainer import CTRTrainer, TrainerConfig

    def job(stop_after=None):
        box = BoxWrapper.set_instance(embedx_dim=4, device=torch.device("cpu"))
        box.initialize_gpu_and_load_model({})
        torch.manual_seed(0)
        tr = CTRTrainer(CtrDnn(num_slots=2, embedx_dim=4, dense_dim=1, hidden=(4,)), box, TrainerConfig())
        seen = []
        for epoch in train_epoch_range(4, save_checkpoint_inter=0, path=str(tmp_path / "job"),
                                       objects=[ModelCheckpoint(trainer=tr, box=box)]):
            seen.append(epoch)
            keys = np.arange(1 + 10 * epoch, 11 + 10 * epoch, dtype=np.uint64)
            rows = box.host_table.fetch_rows(keys, create=True)
            rows[:, 0] = epoch + 1.0
            box.host_table.store_rows(keys, rows)
            with torch.no_grad():
                for p in tr.model.parameters():
                    p.add_(1.0)
            tr.store.mark_dirty() if hasattr(tr.store, "mark_dirty") else None
            if stop_after is not None and epoch == stop_after:
                break                                                        # "crash" before this epoch is saved
        first = next(tr.model.parameters()).detach().clone()
        n = box.host_table.size()
        BoxWrapper.reset_instance()
        return seen, first, n

    seen, w1, n1 = job(stop_after=2)
    assert seen == [0, 1, 2]
    seen, w2, n2 = job()                                                     # epochs 0 and 1 were saved: 2 and 3 remain
    assert seen == [2, 3] and n2 == 40
    torch.testing.assert_close(w2, w1 + 1.0)                                 # 0,1 restored (+2), then 2,3 (+2) vs 0,1,2 (+3)
    seen, w3, n3 = job()
    assert seen == [] and n3 == 40
    torch.testing.assert_close(w3, w2)
    assert CheckpointSaver().get_checkpoint_no(str(tmp_path / "job")) == [3]   # keep=1


def test_without_a_path_the_loop_is_a_range(monkeypatch):
    monkeypatch.delenv("PADDLE_EDL_HDFS_CHECKPOINT_PATH", raising=False)
    assert list(train_epoch_range(3)) == [0, 1, 2]
