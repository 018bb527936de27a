#!/usr/bin/env python
"""Static check for names that are read but defined nowhere (module scope, enclosing scopes, builtins) -- the class of mistake that
only shows up when a rarely taken branch runs, e.g. a GPU-only path on a box without a GPU.  No third-party linter is needed:
Python's own `symtable` says which names a function treats as globals; those must be bound at module level or be builtins.

    python tools/check_names.py [paths...]        # default: the package, bench, benchmarks, baseline harness, experiments, tools
"""
import ast
import builtins
import os
import symtable
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT = ["dgraph_b200", "DGraph", "bench.py", "__graft_entry__.py", "benchmarks", "baseline/ref_bench.py", "baseline/ref_workloads.py",
           "experiments", "tools", "tests"]


def module_bindings(tree: ast.Module):
    names, star = set(), False
    for node in ast.walk(tree):
        if isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
            names.add(node.name)
        elif isinstance(node, ast.Import):
            names.update((a.asname or a.name).split(".")[0] for a in node.names)
        elif isinstance(node, ast.ImportFrom):
            for a in node.names:
                if a.name == "*":
                    star = True
                else:
                    names.add(a.asname or a.name)
        elif isinstance(node, ast.Name) and isinstance(node.ctx, (ast.Store, ast.Del)):
            names.add(node.id)
        elif isinstance(node, ast.Global):
            names.update(node.names)
        elif isinstance(node, ast.ExceptHandler) and node.name:
            names.add(node.name)
    return names, star


def check(path: str):
    src = open(path).read()
    try:
        tree = ast.parse(src, path)
        top = symtable.symtable(src, path, "exec")
    except SyntaxError as e:
        return [f"{path}:{e.lineno}: syntax error: {e.msg}"]
    bound, star = module_bindings(tree)
    if star:
        return []
    known = bound | set(dir(builtins)) | {"__file__", "__name__", "__doc__", "__path__", "__spec__", "__class__", "__builtins__"}
    problems = []

    def visit(tab):
        for sym in tab.get_symbols():
            if sym.is_referenced() and sym.is_global() and sym.get_name() not in known:
                problems.append(f"{path}:{tab.get_lineno()}: '{sym.get_name()}' is read in {tab.get_type()} '{tab.get_name()}' "
                                "but bound nowhere")
        for child in tab.get_children():
            visit(child)

    visit(top)
    return problems


def main(argv):
    paths = argv or DEFAULT
    files = []
    for p in paths:
        p = p if os.path.isabs(p) else os.path.join(ROOT, p)
        if os.path.isdir(p):
            for d, _, fs in os.walk(p):
                if "_ref" in d or "__pycache__" in d:
                    continue
                files += [os.path.join(d, f) for f in fs if f.endswith(".py")]
        elif os.path.exists(p):
            files.append(p)
    problems = [m for f in sorted(files) for m in check(f)]
    print("\n".join(problems) if problems else f"{len(files)} files: every name that is read is bound somewhere")
    return 1 if problems else 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
