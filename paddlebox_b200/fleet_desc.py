"""The sparse-table part of a pslib fleet description (``ps.proto`` text format) -> this framework's sparse config.

A PaddleBox / GPUPS job carries its parameter-server set-up as a protobuf text message (``fleet_desc``:
``server_param { downpour_server_param { downpour_table_param { accessor { ... } } } }``);
``PSGPUTrainer::InitializeGPUServer`` (ps_gpu_trainer.cc:137-254, ``add_sparse_optimizer`` :100-136) reads the
first sparse table's accessor out of it and hands ``PSGPUWrapper`` a flat name -> float map.  This module does the
same without protobuf: a small parser of the text format (nested messages, scalars, repeated fields) and the
accessor-class-specific mapping onto ``SparseOptConfig`` names (+ the optimizer kind).

    conf, optimizer, extra = sparse_config_from_fleet_desc(open("fleet_desc.prototxt").read())
    PSGPUWrapper.get_instance().init_gpu_server({**conf, "optimizer": optimizer})
    # or simply:  PSGPUWrapper.get_instance().init_gpu_server(fleet_desc_text)
"""
from __future__ import annotations

import re

_TOKEN = re.compile(r'\s*(?:(#[^\n]*)|([A-Za-z_][A-Za-z_0-9.]*)|("(?:[^"\\]|\\.)*"|\'(?:[^\'\\]|\\.)*\')|([{}<>:\[\],;])|'
                    r'([-+]?(?:\d+\.?\d*(?:[eE][-+]?\d+)?|\.\d+(?:[eE][-+]?\d+)?)[fF]?|[-+]?inf|nan))')


def _tokens(text: str):
    pos, n = 0, len(text)
    while pos < n:
        m = _TOKEN.match(text, pos)
        if not m:
            if text[pos:].strip() == "":
                return
            raise ValueError(f"fleet desc: cannot read {text[pos:pos + 30]!r}")
        pos = m.end()
        if m.group(1):
            continue
        if m.group(2):
            yield ("name", m.group(2))
        elif m.group(3):
            yield ("str", bytes(m.group(3)[1:-1], "utf-8").decode("unicode_escape"))
        elif m.group(4):
            yield ("sym", m.group(4))
        else:
            v = m.group(5).rstrip("fF")
            yield ("num", float(v) if any(c in v for c in ".eEn") and not v.lstrip("+-").isdigit() else int(v))


def parse_text_format(text: str) -> dict:
    """Protobuf text format -> nested dicts; a field that occurs more than once (or is written as a list) becomes a
    list.  Enum names and ``true`` / ``false`` stay strings / bools."""
    toks = list(_tokens(text))
    i = 0

    def add(msg: dict, key: str, val) -> None:
        if key in msg:
            if not isinstance(msg[key], list):
                msg[key] = [msg[key]]
            msg[key].append(val)
        else:
            msg[key] = val

    def scalar():
        nonlocal i
        kind, v = toks[i]
        i += 1
        if kind == "name":
            return {"true": True, "false": False}.get(v, v)
        if kind in ("num", "str"):
            return v
        raise ValueError(f"fleet desc: unexpected {v!r}")

    def message(close: str | None) -> dict:
        nonlocal i
        msg: dict = {}
        while i < len(toks):
            kind, v = toks[i]
            if kind == "sym" and v in ("}", ">"):
                if close is None:
                    raise ValueError("fleet desc: unbalanced closing brace")
                i += 1
                return msg
            if kind == "sym" and v in (",", ";"):
                i += 1
                continue
            if kind != "name":
                raise ValueError(f"fleet desc: field name expected, got {v!r}")
            key = v
            i += 1
            if i < len(toks) and toks[i] == ("sym", ":"):
                i += 1
            if i >= len(toks):
                raise ValueError(f"fleet desc: value of {key} missing")
            kind, v = toks[i]
            if kind == "sym" and v in ("{", "<"):
                i += 1
                add(msg, key, message(v))
            elif kind == "sym" and v == "[":
                i += 1
                items = []
                while toks[i] != ("sym", "]"):
                    if toks[i] == ("sym", ","):
                        i += 1
                        continue
                    if toks[i][0] == "sym" and toks[i][1] in ("{", "<"):
                        i += 1
                        items.append(message("{"))
                    else:
                        items.append(scalar())
                i += 1
                for it in items:
                    add(msg, key, it)
                if len(items) == 1 and not isinstance(msg[key], list):
                    msg[key] = [msg[key]]
            else:
                add(msg, key, scalar())
        if close is not None:
            raise ValueError("fleet desc: message not closed")
        return msg

    return message(None)


def _first(v):
    return v[0] if isinstance(v, list) else v


def _bounds(sgd: dict, conf: dict, prefix: str) -> None:
    wb = sgd.get("weight_bounds")
    if isinstance(wb, list) and len(wb) == 2:
        conf[prefix + "min_bound"], conf[prefix + "max_bound"] = float(wb[0]), float(wb[1])


def _common_sgd(rule: dict, conf: dict, prefix: str = "") -> str:
    """``SparseCommonSGDRuleParameter`` -> config entries; returns the optimizer kind."""
    name = rule.get("name", "adagrad")
    body = _first(rule.get("adagrad" if name == "std_adagrad" else name, {})) or {}
    if "learning_rate" in body:
        conf[prefix + "learning_rate"] = float(body["learning_rate"])
    if "initial_range" in body:
        conf[prefix + "initial_range"] = float(body["initial_range"])
    if name in ("adagrad", "std_adagrad") and "initial_g2sum" in body:
        conf[prefix + "initial_g2sum"] = float(body["initial_g2sum"])
    if name == "adam":
        for src, dst in (("beta1_decay_rate", "beta1_decay_rate"), ("beta2_decay_rate", "beta2_decay_rate"),
                         ("ada_epsilon", "ada_epsilon")):
            if src in body:
                conf[prefix + dst] = float(body[src])
    _bounds(body, conf, prefix)
    return {"naive": "adagrad", "std_adagrad": "adagrad"}.get(name, name)


def sparse_config_from_fleet_desc(text: str | dict, table: int = 0):
    """``(SparseOptConfig keyword dict, optimizer kind, {"embedx_dim": d} when the accessor names one)`` of sparse
    table ``table`` of a fleet description."""
    desc = parse_text_format(text) if isinstance(text, str) else text
    try:
        tables = desc["server_param"]["downpour_server_param"]["downpour_table_param"]
    except (KeyError, TypeError):
        raise ValueError("fleet desc has no server_param.downpour_server_param.downpour_table_param") from None
    tables = tables if isinstance(tables, list) else [tables]
    sparse = [t for t in tables if str(t.get("type", "PS_SPARSE_TABLE")) == "PS_SPARSE_TABLE"] or tables
    acc = sparse[table].get("accessor", {})
    cls = acc.get("accessor_class", "")
    dp = acc.get("downpour_accessor_param", {}) or {}
    conf: dict = {}
    optimizer = "adagrad"
    if cls in ("DownpourFeatureValueAccessor", "DownpourCtrAccessor", "DownpourCtrDoubleAccessor"):
        sgd = acc.get("sparse_sgd_param", {}) or {}
        conf["nonclk_coeff"] = float(dp.get("nonclk_coeff", 0.1))
        conf["clk_coeff"] = float(dp.get("click_coeff", 1.0))
        for k in ("learning_rate", "initial_g2sum", "initial_range"):
            if k in sgd:
                conf[k] = float(sgd[k])
        _bounds(sgd, conf, "")
        if "embedx_threshold" in acc:
            conf["mf_create_thresholds"] = float(acc["embedx_threshold"])
        for k in ("learning_rate", "initial_g2sum", "initial_range", "min_bound", "max_bound"):
            if k in conf:
                conf["mf_" + k] = conf[k]                     # embedx uses the same rule (ps_gpu_trainer.cc:165-170)
    elif cls == "DownpourSparseValueAccessor":
        optimizer = _common_sgd(acc.get("sparse_commonsgd_param", {}) or {}, conf)
    elif cls in ("DownpourUnitAccessor", "DownpourDoubleUnitAccessor"):
        conf["nonclk_coeff"] = float(dp.get("nonclk_coeff", 0.1))
        conf["clk_coeff"] = float(dp.get("click_coeff", 1.0))
        if "embedx_threshold" in acc:
            conf["mf_create_thresholds"] = float(acc["embedx_threshold"])
        optimizer = _common_sgd(acc.get("embed_sgd_param", {}) or {}, conf)
        _common_sgd(acc.get("embedx_sgd_param", {}) or {}, conf, "mf_")
    else:
        raise ValueError(f"fleet desc: accessor_class {cls!r} is not one the GPU parameter server reads")
    dim = acc.get("embedx_dim")
    extra = {"embedx_dim": int(dim)} if dim is not None else {}
    return conf, optimizer, extra


# keys a job's strategy may carry for the sparse table (pslib/node.py, distributed_strategy.py ``fleet_desc_configs``);
# each may also be prefixed ``embed_`` / ``embedx_`` for the two halves of a unit accessor
STRATEGY_KEYS = frozenset((
    "sparse_table_class", "sparse_compress_in_save", "sparse_shard_num", "sparse_accessor_class", "sparse_learning_rate",
    "sparse_initial_g2sum", "sparse_initial_range", "sparse_weight_bounds", "sparse_fea_dim", "sparse_embedx_dim",
    "sparse_embedx_threshold", "sparse_nonclk_coeff", "sparse_click_coeff", "sparse_base_threshold", "sparse_delta_threshold",
    "sparse_delta_keep_days", "sparse_delete_after_unseen_days", "sparse_show_click_decay_rate", "sparse_delete_threshold",
    "sparse_converter", "sparse_deconverter", "sparse_enable_cache", "sparse_cache_rate", "sparse_cache_file_num",
    "sparse_beta1_decay_rate", "sparse_beta2_decay_rate", "sparse_ada_epsilon", "sparse_optimizer",
    "sparse_ssd_unseenday_threshold", "feature_learning_rate", "nodeid_slot"))


def sparse_config_from_strategy(strategy: dict, prefix: str = ""):
    """The ``strategy`` dict a pslib job gives ``fleet.distributed_optimizer(optimizer, strategy)`` (the reference
    turns it into the fleet description, pslib/node.py:80-330) -> ``(SparseOptConfig keywords, optimizer kind,
    {"embedx_dim": d}, table policy)``.  The table policy holds what the day-end calls use: ``base_threshold`` /
    ``delta_threshold`` (which rows a serving base / delta contains), ``delete_threshold`` /
    ``delete_after_unseen_days`` / ``show_click_decay_rate`` (shrink), ``ssd_unseenday_threshold`` (spill).
    Defaults are the reference's."""
    g = lambda k, d: strategy.get(prefix + k, d)        # noqa: E731
    cls = g("sparse_accessor_class", "DownpourCtrAccessor")
    wb = g("sparse_weight_bounds", None) or [-10, 10]
    conf = {"nonclk_coeff": float(g("sparse_nonclk_coeff", 0.1)), "clk_coeff": float(g("sparse_click_coeff", 1)),
            "mf_create_thresholds": float(g("sparse_embedx_threshold", 10))}
    optimizer = "adagrad"
    if cls in ("DownpourFeatureValueAccessor", "DownpourCtrAccessor", "DownpourCtrDymfAccessor",
               "DownpourCtrDoubleAccessor"):
        rule = {"learning_rate": float(g("sparse_learning_rate", 0.05)), "initial_g2sum": float(g("sparse_initial_g2sum", 3)),
                "initial_range": float(g("sparse_initial_range", 1e-4)), "min_bound": float(wb[0]), "max_bound": float(wb[1])}
        conf.update(rule)
        conf.update({"mf_" + k: v for k, v in rule.items()})
    elif cls in ("DownpourSparseValueAccessor", "DownpourUnitAccessor", "DownpourDoubleUnitAccessor"):
        for pre, key_pre in (("", "embed_"), ("mf_", "embedx_")):
            name = strategy.get(prefix + key_pre + "sparse_optimizer", g("sparse_optimizer", "adagrad"))
            lr_default = 0.001 if name == "adam" else 0.05
            sub = lambda k, d: strategy.get(prefix + key_pre + k, g(k, d))      # noqa: E731
            b = sub("sparse_weight_bounds", None) or [-10, 10]
            conf[pre + "learning_rate"] = float(sub("sparse_learning_rate", lr_default))
            conf[pre + "initial_range"] = float(sub("sparse_initial_range", 1e-4))
            conf[pre + "min_bound"], conf[pre + "max_bound"] = float(b[0]), float(b[1])
            if name in ("adagrad", "std_adagrad"):
                conf[pre + "initial_g2sum"] = float(sub("sparse_initial_g2sum", 3))
            if name == "adam":
                conf[pre + "beta1_decay_rate"] = float(sub("sparse_beta1_decay_rate", 0.9))
                conf[pre + "beta2_decay_rate"] = float(sub("sparse_beta2_decay_rate", 0.999))
                conf[pre + "ada_epsilon"] = float(sub("sparse_ada_epsilon", 1e-8))
            if pre == "":
                optimizer = {"naive": "adagrad", "std_adagrad": "adagrad"}.get(name, name)
    else:
        raise ValueError(f"strategy: sparse_accessor_class {cls!r} is not supported")
    policy = {"base_threshold": float(g("sparse_base_threshold", 1.5)), "delta_threshold": float(g("sparse_delta_threshold", 0.25)),
              "delta_keep_days": float(g("sparse_delta_keep_days", 16)),
              "delete_after_unseen_days": float(g("sparse_delete_after_unseen_days", 30)),
              "ssd_unseenday_threshold": float(g("sparse_ssd_unseenday_threshold", 1)),
              "show_click_decay_rate": float(g("sparse_show_click_decay_rate", 0.98)),
              "delete_threshold": float(g("sparse_delete_threshold", 0.8))}
    return conf, optimizer, {"embedx_dim": int(g("sparse_embedx_dim", 8))}, policy


def format_text(msg: dict, indent: int = 0, enums: tuple = ()) -> str:
    """Nested dicts -> protobuf text (the inverse of ``parse_text_format``): a list is a repeated field, a dict a
    sub-message, a bool ``true`` / ``false``, a str a quoted string unless its field name is listed in ``enums``."""
    pad = "  " * indent
    out = []

    def one(key, v):
        if isinstance(v, dict):
            out.append(f"{pad}{key} {{")
            body = format_text(v, indent + 1, enums)
            if body:
                out.append(body)
            out.append(f"{pad}}}")
        elif isinstance(v, bool):
            out.append(f"{pad}{key}: {'true' if v else 'false'}")
        elif isinstance(v, str):
            if key in enums:
                out.append(f"{pad}{key}: {v}")
            else:
                out.append(pad + key + ': "' + v.replace("\\", "\\\\").replace('"', '\\"').replace("\n", "\\n") + '"')
        else:
            out.append(f"{pad}{key}: {v}")

    for key, val in msg.items():
        if val is None:
            continue
        for v in (val if isinstance(val, (list, tuple)) else [val]):
            one(key, v)
    return "\n".join(out)
