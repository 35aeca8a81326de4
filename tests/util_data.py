"""Synthetic slot-data files in the reference text format (tests and examples)."""
import os
import random

from paddlebox_b200.data_generator import format_instance, make_logkey
from paddlebox_b200.dataset import dense_slot, sparse_slot


def slot_vars(n_sparse=26, n_dense=13, with_label=True):
    """label first (dense 1), then dense features (one dense slot of n_dense), then sparse slots."""
    vs = []
    if with_label:
        vs.append(dense_slot("label", 1))
    if n_dense:
        vs.append(dense_slot("dense", n_dense))
    vs += [sparse_slot(f"slot{i}") for i in range(n_sparse)]
    return vs


def write_files(tmpdir, n_files=2, lines_per_file=200, n_sparse=26, n_dense=13, key_space=5000, max_len=3, seed=0,
                ins_id=False, logkey=False, signal=True):
    """Clicks depend on the ids of slot0, so a model can learn something (AUC > 0.5)."""
    rng = random.Random(seed)
    files = []
    hot = {k: rng.random() for k in range(1, key_space + 1)}
    uid = 0
    for f in range(n_files):
        path = os.path.join(tmpdir, f"part-{f:05d}")
        with open(path, "w") as fp:
            for i in range(lines_per_file):
                slots = []
                sparse = []
                for s in range(n_sparse):
                    n = rng.randint(1, max_len)
                    sparse.append([rng.randint(1, key_space) + s * key_space for _ in range(n)])
                p = hot[sparse[0][0]] if signal else 0.3
                label = 1.0 if rng.random() < p else 0.0
                slots.append([label])
                if n_dense:
                    slots.append([round(rng.random(), 4) + 0.001 for _ in range(n_dense)])
                slots += sparse
                uid += 1
                line = format_instance(slots, ins_id=f"ins{uid}" if ins_id else None,
                                       logkey=make_logkey(uid // 3, 222, uid % 3) if logkey else None)
                fp.write(line + "\n")
        files.append(path)
    return files
