"""Static checks of the Python layer.  Its CUDA-only branches never execute on a CPU box, so two classes of
mistakes that only a GPU run would show are caught here instead: a name that is bound nowhere, and a call
into a native library of an entry point that library does not export."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_no_undefined_names_in_the_python_layer():
    sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
    import undefined_names
    problems = []
    for top in ("paddlebox_b200", "bench.py", "baseline", "benchmarks", "examples", "__graft_entry__.py", "tests"):
        p = os.path.join(ROOT, top)
        files = [p] if p.endswith(".py") else [os.path.join(d, f) for d, _, fs in os.walk(p) for f in fs
                                               if f.endswith(".py") and "_ref" not in d]
        for f in files:
            problems += undefined_names.check(f)
    assert not problems, "\n".join(problems)


def _exports(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True, check=True).stdout
    return {ln.split()[-1] for ln in out.splitlines() if ln.strip()}


def test_every_native_entry_point_called_from_python_is_exported():
    from paddlebox_b200 import _build
    libs = _build.build_all(force=False, verbose=False)
    exported = set()
    for lib in libs.values():
        exported |= _exports(str(lib))
    if "cuda" not in libs:
        import pytest
        pytest.skip("no nvcc: libpbx_cuda.so not built here")
    called = {}
    pat = re.compile(r"\.\s*(pb[xh]_[a-z0-9_]+)\s*\(")
    for d, _, fs in os.walk(os.path.join(ROOT, "paddlebox_b200")):
        for f in fs:
            if f.endswith(".py"):
                src = open(os.path.join(d, f)).read()
                for m in pat.finditer(src):
                    called.setdefault(m.group(1), f)
    for extra in ("bench.py",):
        for m in pat.finditer(open(os.path.join(ROOT, extra)).read()):
            called.setdefault(m.group(1), extra)
    assert len(called) > 80                                   # the pattern really finds the call sites
    missing = {n: f for n, f in called.items() if n not in exported}
    assert not missing, f"called from Python but not exported by libpbx_cuda.so / libpbx_host.so: {missing}"


def _prototypes():
    """name -> parameter count, from the two C headers."""
    protos = {}
    for h in ("include/pbx_cuda.h", "include/pbx_host.h"):
        src = open(os.path.join(ROOT, h)).read()
        src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
        src = re.sub(r"//[^\n]*", " ", src)
        for m in re.finditer(r"\b(pb[xh]_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
            args = m.group(2).strip()
            n = 0 if args in ("", "void") else args.count(",") + 1
            protos[m.group(1)] = n
    return protos


def test_native_calls_pass_as_many_arguments_as_the_prototype_declares():
    """ctypes does not check the argument count of a foreign call: one argument short and the callee reads
    garbage.  Every ``lib.pbx_*(...)`` / ``lib.pbh_*(...)`` call in the package is compared with the header."""
    import ast
    protos = _prototypes()
    assert len(protos) > 150
    bad, seen = [], 0
    files = [os.path.join(d, f) for d, _, fs in os.walk(os.path.join(ROOT, "paddlebox_b200")) for f in fs
             if f.endswith(".py")] + [os.path.join(ROOT, "bench.py")]
    for path in files:
        tree = ast.parse(open(path).read(), path)
        for node in ast.walk(tree):
            if isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and \
                    re.fullmatch(r"pb[xh]_[a-z0-9_]+", node.func.attr) and node.func.attr in protos:
                if any(isinstance(a, ast.Starred) for a in node.args) or node.keywords:
                    continue
                seen += 1
                if len(node.args) != protos[node.func.attr]:
                    bad.append(f"{os.path.relpath(path, ROOT)}:{node.lineno}: {node.func.attr} called with "
                               f"{len(node.args)} arguments, declared with {protos[node.func.attr]}")
    assert seen > 150, seen
    assert not bad, "\n".join(bad)


def test_float_and_64_bit_scalars_are_wrapped_for_ctypes():
    """No ``argtypes`` are declared, so ctypes converts a bare Python int to a 32-bit C int and refuses a bare
    Python float: every ``float`` / ``double`` / 64-bit scalar parameter must be passed through ``C.c_*``."""
    import ast
    params = {}
    for h in ("include/pbx_cuda.h", "include/pbx_host.h"):
        src = open(os.path.join(ROOT, h)).read()
        src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
        src = re.sub(r"//[^\n]*", " ", src)
        for m in re.finditer(r"\b(pb[xh]_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
            a = m.group(2).strip()
            params[m.group(1)] = [] if a in ("", "void") else [x.strip() for x in a.split(",")]
    bad, seen = [], 0
    files = [os.path.join(d, f) for d, _, fs in os.walk(os.path.join(ROOT, "paddlebox_b200")) for f in fs
             if f.endswith(".py")] + [os.path.join(ROOT, "bench.py")]
    for path in files:
        for node in ast.walk(ast.parse(open(path).read(), path)):
            if not (isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and node.func.attr in params):
                continue
            ps = params[node.func.attr]
            if len(ps) != len(node.args):
                continue
            for p, a in zip(ps, node.args):
                if "*" in p:
                    continue
                src = ast.unparse(a)
                if re.search(r"\b(float|double)\b", p):
                    seen += 1
                    if not re.match(r"C\.c_(float|double)\(", src):
                        bad.append(f"{os.path.relpath(path, ROOT)}:{node.lineno}: {node.func.attr}({p}) <- {src}")
                elif re.search(r"\b(u?int64_t|size_t|long long)\b", p):
                    seen += 1
                    if not re.match(r"C\.c_(u?int64|size_t|u?longlong)\(", src):
                        bad.append(f"{os.path.relpath(path, ROOT)}:{node.lineno}: {node.func.attr}({p}) <- {src}")
    assert seen > 100, seen
    assert not bad, "\n".join(bad)


def test_no_method_is_defined_twice_in_a_class():
    """A second ``def`` of the same name silently replaces the first (it did: ``BoxWrapper.get_lr_map``)."""
    import ast
    dup = []
    for d, _, fs in os.walk(os.path.join(ROOT, "paddlebox_b200")):
        for f in fs:
            if not f.endswith(".py"):
                continue
            path = os.path.join(d, f)
            for node in ast.walk(ast.parse(open(path).read(), path)):
                if not isinstance(node, ast.ClassDef):
                    continue
                seen = {}
                for b in node.body:
                    if isinstance(b, (ast.FunctionDef, ast.AsyncFunctionDef)):
                        deco = [ast.unparse(x) for x in b.decorator_list]
                        if any(k in x for x in deco for k in ("setter", "property", "overload")):
                            continue
                        if b.name in seen:
                            dup.append(f"{os.path.relpath(path, ROOT)}: {node.name}.{b.name} at lines {seen[b.name]} and {b.lineno}")
                        seen[b.name] = b.lineno
    assert not dup, "\n".join(dup)


def test_every_flag_of_the_reference_paddlebox_sources_exists():
    """The FLAGS_* the reference's PaddleBox sources define or read (collected once from boxps worker / trainer,
    data_set, data_feed, box_wrapper, ps_gpu_wrapper, heter_ps, metrics and the PaddleBox block of
    platform/flags.cc) all exist here — wired, or accepted with the reason in their help text."""
    from paddlebox_b200.flags import FLAGS
    names = """padbox_record_pool_max_size padbox_slotpool_thread_num padbox_dataset_shuffle_thread_num
    padbox_dataset_merge_thread_num padbox_dataset_disable_shuffle padbox_dataset_disable_polling
    padbox_dataset_enable_unrollinstance padbox_auc_runner_mode padbox_dump_debug_lineid padbox_slotrecord_extend_dim
    enable_pullpush_dedup_keys enable_pull_box_padding_zero use_gpu_replica_cache gpu_replica_cache_dim
    enable_dense_nccl_barrier enable_sync_dense_moment enable_shuffle_by_searchid enable_binding_train_cpu
    enable_ins_parser_file enable_ins_parser_add_file_path enable_dump_main_program enable_force_hbm_recyle
    enable_force_mem_recyle enbale_slotpool_auto_clear enable_slotpool_wait_release enable_slotrecord_reset_shrink
    enable_print_dump_field_debug enable_print_dump_info_debug dump_filed_same_as_aibox lineid_have_extend_info
    enable_opt_get_features gpugraph_load_node_list_into_hbm graph_get_neighbor_id check_nan_inf""".split()
    missing = [n for n in names if not hasattr(FLAGS, n)]
    assert not missing, missing
