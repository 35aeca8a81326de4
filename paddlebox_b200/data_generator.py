"""Text instance format writers (reference ``MultiSlotDataGenerator``,
python/paddle/distributed/fleet/data_generator/data_generator.py:284-370).

One instance per line: for every slot, in slot order, ``<n> <v_1> ... <v_n>`` with n >= 1;
optionally preceded by ``1 <ins_id>`` and ``1 <logkey>`` groups."""
from __future__ import annotations

import sys
from typing import Iterable, Sequence


class DataGenerator:
    def __init__(self):
        self._proto_info = None
        self.batch_size_ = 32

    def set_batch(self, batch_size: int):
        self.batch_size_ = batch_size

    def generate_sample(self, line):
        """Override: return a generator yielding ``[(slot_name, [values...]), ...]``."""
        raise NotImplementedError

    def generate_batch(self, samples):
        def local_iter():
            for s in samples:
                yield s
        return local_iter

    def _gen_str(self, line) -> str:
        raise NotImplementedError

    def run_from_stdin(self):
        self.run_from_iterable(sys.stdin, sys.stdout)

    def run_from_memory(self):
        batch = []
        for sample in self.generate_sample(None)():
            batch.append(sample)
            if len(batch) == self.batch_size_:
                for s in self.generate_batch(batch)():
                    sys.stdout.write(self._gen_str(s))
                batch = []
        for s in self.generate_batch(batch)():
            sys.stdout.write(self._gen_str(s))

    def run_from_iterable(self, lines: Iterable[str], out):
        batch = []
        for line in lines:
            for sample in self.generate_sample(line)():
                if sample is None:
                    continue
                batch.append(sample)
                if len(batch) == self.batch_size_:
                    for s in self.generate_batch(batch)():
                        out.write(self._gen_str(s))
                    batch = []
        for s in self.generate_batch(batch)():
            out.write(self._gen_str(s))


class MultiSlotDataGenerator(DataGenerator):
    def _gen_str(self, line) -> str:
        if not isinstance(line, (list, tuple)):
            raise ValueError("the output of generate_sample must be a list or tuple of (name, values) pairs")
        parts = []
        if self._proto_info is None:
            self._proto_info = []
            for name, elements in line:
                if not elements:
                    raise ValueError(f"slot {name!r} has no value; every slot needs at least one (pad with 0)")
                self._proto_info.append((name, "float" if isinstance(elements[0], float) else "uint64"))
        elif len(line) != len(self._proto_info):
            raise ValueError("the slot count of this instance differs from the first instance")
        for i, (name, elements) in enumerate(line):
            if name != self._proto_info[i][0]:
                raise ValueError(f"slot order changed: expected {self._proto_info[i][0]!r}, got {name!r}")
            if not elements:
                raise ValueError(f"slot {name!r} has no value")
            if self._proto_info[i][1] == "uint64" and isinstance(elements[0], float):
                self._proto_info[i] = (name, "float")
            parts.append(str(len(elements)))
            parts.extend(str(e) for e in elements)
        return " ".join(parts) + "\n"


class MultiSlotStringDataGenerator(MultiSlotDataGenerator):
    def _gen_str(self, line) -> str:
        parts = []
        for name, elements in line:
            parts.append(str(len(elements)))
            parts.extend(elements)
        return " ".join(parts) + "\n"


def format_instance(slots: Sequence[Sequence], ins_id: str | None = None, logkey: str | None = None) -> str:
    """Format one instance from per-slot value lists (no generator class needed)."""
    parts = []
    if ins_id is not None:
        parts += ["1", ins_id]
    if logkey is not None:
        parts += ["1", logkey]
    for vals in slots:
        if len(vals) == 0:
            raise ValueError("every slot needs at least one value")
        parts.append(str(len(vals)))
        parts.extend(repr(float(v)) if isinstance(v, float) else str(int(v)) for v in vals)
    return " ".join(parts)


def make_logkey(search_id: int, cmatch: int, rank: int) -> str:
    """Hex logkey with cmatch at [11,14), rank at [14,16), search_id at [16,32)
    (parser_log_key, data_feed.cc:2385-2396)."""
    return "0" * 11 + f"{cmatch:03x}" + f"{rank:02x}" + f"{search_id:016x}"
