"""Global (cross-rank) record shuffle.

Reference: ``PadBoxSlotDataset::ShuffleData`` hash-partitions the loaded records to ranks — by
``search_id`` (``FLAGS_enable_shuffle_by_searchid``), by XXH64 of the first 32 bytes of the
instance id, or randomly — serialises them with ``BinaryArchive`` and ships them through the
closed ``PaddleShuffler`` RPC; the receiver de-serialises into its own record list
(framework/data_set.cc:2422-2604).  Here every rank serialises one binary archive per destination
(``csrc/host/dataset.cc::save_partitioned``; the record wire format is the dataset archive format)
and the archives travel by one of two transports:

* ``"network"`` — the archives are written to rank-private scratch files, their bytes exchanged with
  ONE ``all_to_all`` over ``torch.distributed`` (gloo / NCCL over the NICs) and loaded by the
  receiver: no file system shared between nodes is needed (the ``PaddleShuffler`` role);
* ``"files"``   — the archives are written into a directory every rank can read (tmpfs on one node,
  a shared file system across nodes); ranks meet at a barrier and load what is addressed to them.
"""
from __future__ import annotations

import glob
import os

MODES = {"random": 0, "search_id": 1, "ins_id": 2}


def _exchange_blobs(blobs, group):
    """``blobs[d]`` goes to rank d; returns the list of blobs received, by source rank."""
    import numpy as np
    import torch
    import torch.distributed as dist
    world = len(blobs)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    send_sizes = torch.tensor([len(b) for b in blobs], dtype=torch.int64, device=dev)
    recv_sizes = torch.empty_like(send_sizes)
    dist.all_to_all_single(recv_sizes, send_sizes, group=group)
    sc, rc = send_sizes.tolist(), recv_sizes.tolist()
    joined = b"".join(blobs)
    send = torch.from_numpy(np.frombuffer(joined, dtype=np.uint8).copy()) if joined else torch.empty(0, dtype=torch.uint8)
    send = send.to(dev)
    recv = torch.empty(sum(rc), dtype=torch.uint8, device=dev)
    dist.all_to_all_single(recv, send, output_split_sizes=rc, input_split_sizes=sc, group=group)
    flat = recv.cpu().numpy().tobytes()
    out, off = [], 0
    for n in rc:
        out.append(flat[off:off + n])
        off += n
    assert len(out) == world
    return out


def global_shuffle_records(dataset, work_dir: str, mode: str = "random", rank: int | None = None,
                           world: int | None = None, group=None, seed: int = 0, tag: str = "shuffle",
                           transport: str = "files") -> int:
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
    nat.set("rank", rank)
    if transport == "network":
        import tempfile
        with tempfile.TemporaryDirectory(prefix=f"pbx_shuffle_r{rank}_") as scratch:     # rank-private
            prefix = os.path.join(scratch, tag)
            nat.save_partitioned(prefix, world, MODES[mode], seed)
            blobs = []
            for dst in range(world):
                with open(f"{prefix}.{rank}.{dst}", "rb") as f:
                    blobs.append(f.read())
                os.remove(f"{prefix}.{rank}.{dst}")
            nat.release_memory()
            for src, blob in enumerate(_exchange_blobs(blobs, group)):
                path = f"{prefix}.in.{src}"
                with open(path, "wb") as f:
                    f.write(blob)
                nat.load_archive(path)
                os.remove(path)
        nat.local_shuffle(seed + rank)
        return nat.num_records()
    os.makedirs(work_dir, exist_ok=True)
    prefix = os.path.join(work_dir, tag)
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
