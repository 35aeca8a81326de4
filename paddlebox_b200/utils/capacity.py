"""HBM / host-memory budget of a configuration — sizing shards for 180 GB per B200 before a job starts.

The numbers restate the allocations of ``pbx_table_create`` (csrc/cuda/table.cu: one key, one value row
and one dirty byte per table slot) and ``pbx_sparse_create`` (csrc/cuda/sparse_a2a.cu: symmetric
mailboxes for keys / pull rows / push rows per (rank, rank) pair, the batch dedup table, the merge
accumulator), plus the dense side (fp32 parameters + gradients in symmetric memory, two Adam moments,
bf16 operand copies of the towers).  Reference role: ``CheckNeedLimitMem`` / the HBM pool sizing the
closed library does internally; here the arithmetic is in the open so that a config can be checked on a
machine without a GPU.
"""
from __future__ import annotations

from dataclasses import dataclass

from .. import _native as N
from ..flags import FLAGS

OPTIMIZERS = {"adagrad": 0, "adam": 1, "shared_adam": 2}
B200_HBM_BYTES = 180 * 10 ** 9


def _al(x: int) -> int:
    return (x + 255) // 256 * 256


@dataclass
class Budget:
    table_bytes: int
    engine_bytes: int
    dense_bytes: int
    host_bytes_per_key: int
    table_capacity: int
    mailbox_cap: int

    @property
    def hbm_bytes(self) -> int:
        return self.table_bytes + self.engine_bytes + self.dense_bytes

    def fits(self, hbm_bytes: int = B200_HBM_BYTES, reserve: float = 0.08) -> bool:
        """True when everything fits with ``reserve`` of the device left for activations, the CUDA context
        and allocator slack."""
        return self.hbm_bytes <= hbm_bytes * (1.0 - reserve)

    def report(self) -> str:
        gb = 1e9
        return (f"table {self.table_bytes / gb:.2f} GB ({self.table_capacity} slots) + engine {self.engine_bytes / gb:.2f} GB "
                f"(mailbox cap {self.mailbox_cap}) + dense {self.dense_bytes / gb:.3f} GB = {self.hbm_bytes / gb:.2f} GB HBM; "
                f"host tier {self.host_bytes_per_key} B per key")


def plan(keys_per_gpu: int, embedx_dim: int, optimizer: str = "adagrad", max_keys_per_batch: int = 1 << 20,
         world_size: int = 1, dense_params: int = 0, load_factor: float | None = None,
         table_capacity: int | None = None) -> Budget:
    """Budget of one GPU holding ``keys_per_gpu`` rows of width ``embedx_dim``."""
    L = N.make_layout(embedx_dim, OPTIMIZERS[optimizer])
    lf = FLAGS.gpugraph_hbm_table_load_factor if load_factor is None else load_factor
    cap_slots = int(table_capacity or -(-keys_per_gpu // lf))
    table = cap_slots * (8 + L.stride * 4 + 1) + 16
    R = world_size
    if R == 1:
        cap = max_keys_per_batch
    else:
        cap = min(max_keys_per_batch, int(FLAGS.pbx_sparse_cap_factor * max_keys_per_batch / R) + 1024)
    dd = 1024
    while dd < max_keys_per_batch * 2:
        dd <<= 1
    symm = _al(64 * 8) + _al(64 * 4) + _al(R * cap * 8) + _al(R * cap * L.pull_stride * 4) + _al(R * cap * L.push_stride * 4)
    engine = symm + dd * 8 + dd * 4 + 2 * 64 * 4 + R * cap * 4 + R * cap * L.push_stride * 4 + 256
    # dense: params + (grads | scratch) in symmetric memory + two moments, all fp32, + bf16 operand copies
    dense = dense_params * (4 + 2 * 4 + 2 * 4 + 2 * 2)
    host_per_key = L.stride * 4 + 8 + 1 + 12           # row + key + dirty byte + index cell
    return Budget(table, engine, dense, host_per_key, cap_slots, cap)
