"""Field / parameter dump (reference dump_fields / dump_param written by a thread pool into
rolling part files, boxps_worker.cc:1595-1855: one line per instance
``lineid\\tfield:len:v1:v2...``; sampling by ``dump_mode`` / ``dump_interval``)."""
from __future__ import annotations

import os
import queue
import threading
import zlib

import torch

from ..flags import FLAGS


class FieldDumper:
    def __init__(self, path: str, fields, dump_mode: int = 0, dump_interval: int = 10, num_threads: int = 1,
                 part_bytes: int = 2 << 30, rank: int = 0, user_filename: str = ""):
        """dump_mode 0 = all instances, 1 = hash(lineid) % interval == 0, 2 = every
        ``dump_interval``-th instance."""
        os.makedirs(path, exist_ok=True)
        self.path, self.fields = path, list(fields)
        self.mode, self.interval = dump_mode, max(1, dump_interval)
        self.part_bytes, self.rank = part_bytes, rank
        self.user_filename = user_filename        # TrainerDesc.user_define_dump_filename: part-<name>-<n>
        self._q: "queue.Queue" = queue.Queue(maxsize=64)
        self._part, self._written, self._fp = 0, 0, None
        self._seen = 0
        self._threads = [threading.Thread(target=self._loop, daemon=True) for _ in range(max(1, num_threads))]
        self._lock = threading.Lock()
        for t in self._threads:
            t.start()

    def _open(self):
        if self._fp:
            self._fp.close()
        stem = self.user_filename if self.user_filename else f"{self.rank:03d}"         # (GetDumpPath, ps_gpu_trainer.cc:256-266)
        self._fp = open(os.path.join(self.path, f"part-{stem}-{self._part:05d}"), "w")
        self._part += 1
        self._written = 0

    def _loop(self):
        while True:
            item = self._q.get()
            if item is None:
                self._q.task_done()
                return
            text = "".join(item)
            with self._lock:
                if self._fp is None or self._written + len(text) > self.part_bytes:
                    self._open()
                self._fp.write(text)
                self._written += len(text)
            self._q.task_done()

    def _keep(self, lineid: str) -> bool:
        if FLAGS.padbox_dump_debug_lineid:                  # debug: only this line id (boxps_worker.cc:47-49)
            return lineid == FLAGS.padbox_dump_debug_lineid
        if self.mode == 0:
            return True
        if self.mode == 1:
            return zlib.crc32(lineid.encode()) % self.interval == 0
        self._seen += 1
        return self._seen % self.interval == 0

    def dump_batch(self, ins_ids, tensors: dict):
        host = {k: tensors[k].detach().float().cpu().reshape(len(ins_ids), -1) for k in self.fields if k in tensors}
        lines = []
        aibox, ext = FLAGS.dump_filed_same_as_aibox, FLAGS.lineid_have_extend_info
        debug = FLAGS.enable_print_dump_field_debug
        for i, lid in enumerate(ins_ids):
            if not self._keep(lid):
                continue
            # FLAGS_lineid_have_extend_info: "id<space>extra" -> the id leads, the extra part closes the line
            head, _, extra = lid.partition(" ") if ext else (lid, "", "")
            parts = [head]
            for k, t in host.items():
                row = t[i].tolist()
                # FLAGS_dump_filed_same_as_aibox: "name:v:v" with the name cut at its first '.', no count
                name = k.split(".", 1)[0] if aibox else f"{k}:{len(row)}"
                parts.append(name + ":" + ":".join(f"{v:.6g}" for v in row))
                if debug:
                    print(f"[dump] lineid=[{lid}] name={k} width={len(row)}", flush=True)
            if ext and extra:
                parts.append(extra)
            lines.append("\t".join(parts) + "\n")
        if FLAGS.enable_print_dump_info_debug:
            print(f"[dump] rank {self.rank}: dump ins count={len(lines)} of {len(ins_ids)}, fields={len(host)}", flush=True)
        if lines:
            self._q.put(lines)

    def dump_params(self, named_params):
        lines = []
        for name, p in named_params:
            flat = p.detach().float().cpu().reshape(-1).tolist()
            lines.append(f"{name}:{len(flat)}:" + ":".join(f"{v:.6g}" for v in flat) + "\n")
        self._q.put(lines)

    def close(self):
        self._q.join()
        for _ in self._threads:
            self._q.put(None)
        for t in self._threads:
            t.join()
        if self._fp:
            self._fp.close()
            self._fp = None
