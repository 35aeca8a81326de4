"""Global (cross-rank) record shuffle.

Reference: ``PadBoxSlotDataset::ShuffleData`` hash-partitions the loaded records to ranks — by
``search_id`` (``FLAGS_enable_shuffle_by_searchid``), by XXH64 of the first 32 bytes of the
instance id, or randomly — serialises them with ``BinaryArchive`` and ships them through the
closed ``PaddleShuffler`` RPC; the receiver de-serialises into its own record list
(framework/data_set.cc:2422-2604).  Here every rank writes one binary archive per destination
(``csrc/host/dataset.cc::save_partitioned``) into a directory all ranks can read (local disk /
tmpfs on one node, a shared file system across nodes), meets at a barrier, and loads the
archives addressed to it.  The record wire format is the dataset archive format.
"""
from __future__ import annotations

import glob
import os

MODES = {"random": 0, "search_id": 1, "ins_id": 2}


def global_shuffle_records(dataset, work_dir: str, mode: str = "random", rank: int | None = None,
                           world: int | None = None, group=None, seed: int = 0, tag: str = "shuffle") -> int:
    """Re-partition ``dataset``'s in-memory records over the ranks of ``group``.
    Returns the number of records this rank holds afterwards."""
    import torch.distributed as dist
    if rank is None:
        rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size(group) if dist.is_initialized() else 1
    nat = dataset._native
    if world == 1:
        nat.local_shuffle(seed)
        return nat.num_records()
    os.makedirs(work_dir, exist_ok=True)
    prefix = os.path.join(work_dir, tag)
    nat.set("rank", rank)
    nat.save_partitioned(prefix, world, MODES[mode], seed)
    dist.barrier(group=group)
    nat.release_memory()
    for src in range(world):
        nat.load_archive(f"{prefix}.{src}.{rank}")
    dist.barrier(group=group)
    for path in glob.glob(f"{prefix}.{rank}.*"):
        os.remove(path)
    nat.local_shuffle(seed + rank)
    return nat.num_records()
