"""``DataFeedDesc``: the feed configuration as an object built from a prototxt file.

Reference: ``python/paddle/fluid/data_feed_desc.py`` (:21-250) over ``framework/data_feed.proto``.  The file is
protobuf text (``name``, ``batch_size``, ``multi_slot_desc { slots {...} }``, ``pipe_command`` ...); it is read
and written without protobuf (``fleet_desc.parse_text_format`` / ``format_text``).  ``Dataset.set_data_feed_desc``
takes ``desc()``.
"""
from __future__ import annotations

from .fleet_desc import format_text, parse_text_format


class DataFeedDesc:
    def __init__(self, proto_file: str):
        with open(proto_file) as f:
            self.proto_desc = parse_text_format(f.read())
        self.proto_desc.setdefault("pipe_command", "cat")
        msd = self.proto_desc.setdefault("multi_slot_desc", {})
        slots = msd.get("slots", [])
        msd["slots"] = slots if isinstance(slots, list) else [slots]
        self._index = {s["name"]: i for i, s in enumerate(msd["slots"])} \
            if self.proto_desc.get("name") == "MultiSlotDataFeed" else None

    def set_batch_size(self, batch_size: int) -> None:
        self.proto_desc["batch_size"] = int(batch_size)

    def _mark(self, names, field: str, what: str) -> None:
        if self._index is None:
            raise ValueError(f"Only MultiSlotDataFeed needs {what}, please check your datafeed.proto")
        for name in names:
            self.proto_desc["multi_slot_desc"]["slots"][self._index[name]][field] = True

    def set_dense_slots(self, dense_slots_name) -> None:
        """The named slots are dense: fed as one fixed-width tensor instead of a ragged one."""
        self._mark(dense_slots_name, "is_dense", "set_dense_slots")

    def set_use_slots(self, use_slots_name) -> None:
        """The named slots are used by the model (a file may carry more slots than a model reads)."""
        self._mark(use_slots_name, "is_used", "set_use_slots")

    def desc(self) -> str:
        return format_text(self.proto_desc)

    __str__ = desc
