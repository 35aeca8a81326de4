"""Model manifests ("donefiles") and resume.

PaddleBox jobs publish every saved model by appending one JSON line to a donefile next to the
models — ``xbox_base_done.txt`` for day-level base models, ``xbox_patch_done.txt`` for per-pass
deltas — and both the online serving side and a restarted training job find "the latest base and
every delta after it" there (the reference leaves this to the user scripts around
``SaveBase`` / ``SaveDelta``, fleet/box_wrapper.cc:1286-1315; its recovery model is restart from
the last saved model).  This module is that convention:

* ``ModelManifest.record_base / record_delta`` append entries (atomically: write + rename),
* ``ModelManifest.latest()`` returns the newest base and the deltas recorded after it,
* ``resume(box, manifest)`` loads them into a ``BoxWrapper`` (training state when the entry names a
  batch model, otherwise the serving models).
"""
from __future__ import annotations

import json
import os
import time

BASE_DONE, PATCH_DONE = "xbox_base_done.txt", "xbox_patch_done.txt"


class ModelManifest:
    def __init__(self, root: str):
        self.root = root
        os.makedirs(root, exist_ok=True)

    # ---- writing -------------------------------------------------------------------------------
    def _append(self, name: str, entry: dict) -> dict:
        path = os.path.join(self.root, name)
        # keep the complete, parsable lines only: a torn tail left by a crashed writer must not swallow the
        # new entry
        old = "".join(json.dumps(e, sort_keys=True) + "\n" for e in self._read(name))
        tmp = f"{path}.tmp.{os.getpid()}"          # (one writer per manifest by convention: rank 0 of the job)
        with open(tmp, "w") as f:
            f.write(old + json.dumps(entry, sort_keys=True) + "\n")
            f.flush()
            os.fsync(f.fileno())
        os.replace(tmp, path)                      # readers never see a half-written line
        return entry

    def record_base(self, day: str, xbox_path: str, batch_model_path: str = "", dense_path: str = "",
                    pass_id: int = 0) -> dict:
        return self._append(BASE_DONE, {"id": str(int(time.time() * 1000)), "key": f"{day}", "day": str(day),
                                        "pass_id": int(pass_id), "input": xbox_path, "batch_model": batch_model_path,
                                        "dense": dense_path, "record_count": "111111", "job_name": "default_job_name"})

    def record_delta(self, day: str, pass_id: int, xbox_path: str, dense_path: str = "") -> dict:
        return self._append(PATCH_DONE, {"id": str(int(time.time() * 1000)), "key": f"{day}/{pass_id}", "day": str(day),
                                         "pass_id": int(pass_id), "input": xbox_path, "dense": dense_path,
                                         "record_count": "111111", "job_name": "default_job_name"})

    # ---- reading ---------------------------------------------------------------------------------
    def _read(self, name: str) -> list:
        path = os.path.join(self.root, name)
        if not os.path.exists(path):
            return []
        out = []
        with open(path) as f:
            for line in f:
                line = line.strip()
                if line:
                    try:
                        out.append(json.loads(line))
                    except json.JSONDecodeError:
                        break                     # a torn last line (crash while a foreign writer appended)
        return out

    def bases(self) -> list:
        return self._read(BASE_DONE)

    def deltas(self) -> list:
        return self._read(PATCH_DONE)

    def latest(self):
        """``(base entry or None, [delta entries recorded after that base, oldest first])``."""
        bases = self.bases()
        base = bases[-1] if bases else None
        deltas = self.deltas()
        if base is not None:
            deltas = [d for d in deltas if int(d["id"]) > int(base["id"])
                      or (d["day"], d["pass_id"]) > (base["day"], base["pass_id"])]
        return base, deltas


def resume(box, manifest: ModelManifest, serving: bool = False, model=None) -> dict:
    """Load the newest base (its batch model when there is one and ``serving`` is False, else the xbox
    base) and every later delta into ``box``; returns what was loaded.  Deltas are serving-format rows:
    weights and show / click are restored, optimizer state of those rows restarts.  With ``model`` the dense
    parameters recorded with the newest entry (``dense`` directory, ``checkpoint.save_persistables`` layout)
    are loaded into it as well."""
    base, deltas = manifest.latest()
    loaded = {"base": None, "deltas": [], "rows": 0}
    if base is not None:
        use_batch = bool(base.get("batch_model")) and not serving and os.path.isdir(base["batch_model"])
        path = base["batch_model"] if use_batch else base["input"]
        loaded["rows"] += box.load_model(path, serving=not use_batch)
        loaded["base"] = path
    for d in deltas:
        loaded["rows"] += box.load_model(d["input"], serving=True)
        loaded["deltas"].append(d["input"])
    if base is not None and base.get("day"):
        box.set_date(base["day"])
    if model is not None:
        newest = (deltas[-1] if deltas else base) or {}
        dense_dir = newest.get("dense") or (base or {}).get("dense")
        if dense_dir and os.path.isdir(dense_dir):
            from ..checkpoint import load_persistables
            load_persistables(model, dense_dir, strict=False)
            fn = getattr(model, "mark_weights_dirty", None)
            if fn:
                fn()
            loaded["dense"] = dense_dir
    return loaded
