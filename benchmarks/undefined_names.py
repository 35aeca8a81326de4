#!/usr/bin/env python
"""Poor man's pyflakes (no linter in the image): names that are read somewhere in a module but bound nowhere in
the enclosing scopes, the module or the builtins.  Its purpose here: the CUDA-only branches of the Python layer do
not execute on a CPU box, so a typo in one would only show on a GPU.

    python benchmarks/undefined_names.py [paths...]        # default: paddlebox_b200 bench.py baseline tests benchmarks
"""
import ast
import builtins
import os
import sys


class Scope:
    def __init__(self, node, parent):
        self.node, self.parent, self.names = node, parent, set()


def bound_names(target, out):
    for n in ast.walk(target):
        if isinstance(n, ast.Name) and isinstance(n.ctx, (ast.Store, ast.Del)):
            out.add(n.id)


def collect(scope_node, scope):
    """Names bound directly in this scope (not in nested function / class scopes)."""
    def visit(n):
        for child in ast.iter_child_nodes(n):
            if isinstance(child, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
                scope.names.add(child.name)
                for d in child.decorator_list:
                    visit(d)
                continue
            if isinstance(child, ast.Lambda):
                continue
            if isinstance(child, (ast.Import, ast.ImportFrom)):
                for a in child.names:
                    scope.names.add((a.asname or a.name).split(".")[0])
            elif isinstance(child, ast.Name) and isinstance(child.ctx, (ast.Store, ast.Del)):
                scope.names.add(child.id)
            elif isinstance(child, ast.ExceptHandler) and child.name:
                scope.names.add(child.name)
            elif isinstance(child, (ast.Global, ast.Nonlocal)):
                scope.names.update(child.names)
            elif isinstance(child, (ast.MatchAs, ast.MatchStar)) and getattr(child, "name", None):
                scope.names.add(child.name)
            visit(child)
    if isinstance(scope_node, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda)):
        a = scope_node.args
        for arg in a.posonlyargs + a.args + a.kwonlyargs + ([a.vararg] if a.vararg else []) + ([a.kwarg] if a.kwarg else []):
            scope.names.add(arg.arg)
    if isinstance(scope_node, ast.Lambda):
        visit(ast.Expr(scope_node.body))
        bound_names(scope_node.body, scope.names)   # walrus
    else:
        visit(scope_node)


def check(path):
    src = open(path).read()
    try:
        tree = ast.parse(src, path)
    except SyntaxError as e:
        return [f"{path}:{e.lineno}: syntax error {e.msg}"]
    problems = []
    builtin = set(dir(builtins)) | {"__file__", "__name__", "__doc__", "__builtins__", "__spec__", "__package__"}

    def walk(node, scope):
        for child in ast.iter_child_nodes(node):
            if isinstance(child, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda)):
                # defaults and decorators evaluate in the enclosing scope
                for d in getattr(child, "decorator_list", []):
                    walk_expr(d, scope)
                for d in child.args.defaults + [k for k in child.args.kw_defaults if k is not None]:
                    walk_expr(d, scope)
                s = Scope(child, scope)
                collect(child, s)
                body = child.body if isinstance(child.body, list) else [child.body]
                for b in body:
                    walk_expr(b, s)
            elif isinstance(child, ast.ClassDef):
                for d in child.decorator_list + child.bases + [k.value for k in child.keywords]:
                    walk_expr(d, scope)
                s = Scope(child, scope)
                collect(child, s)
                for b in child.body:
                    walk_expr(b, s)
            else:
                walk_expr(child, scope)

    def walk_expr(node, scope):
        if isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda, ast.ClassDef)):
            walk(ast.Module(body=[node], type_ignores=[]) if not isinstance(node, ast.Lambda) else ast.Expr(node), scope)
            return
        if isinstance(node, (ast.ListComp, ast.SetComp, ast.DictComp, ast.GeneratorExp)):
            s = Scope(node, scope)
            for g in node.generators:
                bound_names(g.target, s.names)
            for sub in ast.iter_child_nodes(node):
                walk_expr(sub, s)
            return
        if isinstance(node, ast.Name) and isinstance(node.ctx, ast.Load):
            s, ok = scope, False
            while s is not None:
                # class scopes are not visible from nested functions, but being lenient there only hides problems
                if node.id in s.names:
                    ok = True
                    break
                s = s.parent
            if not ok and node.id not in builtin:
                problems.append(f"{path}:{node.lineno}: undefined name '{node.id}'")
        walk(node, scope)

    top = Scope(tree, None)
    collect(tree, top)
    walk(tree, top)
    return problems


def main():
    roots = sys.argv[1:] or ["paddlebox_b200", "bench.py", "baseline", "tests", "benchmarks", "examples", "__graft_entry__.py"]
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    files = []
    for r in roots:
        r = os.path.join(here, r) if not os.path.isabs(r) else r
        if os.path.isdir(r):
            for d, _, fs in os.walk(r):
                if "_ref" in d or "__pycache__" in d:
                    continue
                files += [os.path.join(d, f) for f in fs if f.endswith(".py")]
        elif os.path.exists(r):
            files.append(r)
    out = []
    for f in sorted(files):
        out += check(f)
    for p in out:
        print(p)
    print(f"{len(files)} files, {len(out)} undefined names")
    return 1 if out else 0


if __name__ == "__main__":
    sys.exit(main())
