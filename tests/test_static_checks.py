"""Static checks that stand in for running GPU-only branches on a box without a GPU: every name a function reads is bound somewhere
(`tools/check_names.py`), and every native entry point the Python layer names exists in the built libraries."""
import glob
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_name_that_is_read_is_bound():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "check_names.py")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-4000:]


def _exports(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True).stdout
    return {line.split()[-1] for line in out.splitlines() if line.strip()}


def test_every_native_symbol_named_in_python_is_exported():
    libs = [os.path.join(ROOT, "dgraph_b200", "lib", n) for n in ("libdgraph_b200.so", "libdgraph_b200_host.so")]
    if not all(os.path.exists(p) for p in libs):
        pytest.skip("libraries not built (python -m dgraph_b200.build)")
    have = set().union(*(_exports(p) for p in libs))
    used = {}
    for f in glob.glob(os.path.join(ROOT, "**", "*.py"), recursive=True):
        rel = os.path.relpath(f, ROOT)
        if rel.startswith(("baseline/_ref", "tests", "gpurun_out", "build")):
            continue
        for m in re.finditer(r"\b(dgbh?_[a-z0-9_]+)\b", open(f).read()):
            used.setdefault(m.group(1), set()).add(rel)
    missing = {k: sorted(v) for k, v in used.items() if k not in have}
    assert not missing, f"named in Python but not exported by the libraries: {missing}"


def _strip_comments(text):
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.sub(r"//[^\n]*", "", text)


def _param_counts(text, terminator):
    out = {}
    for m in re.finditer(r"\b(dgb_[a-z0-9_]+)\s*\(([^;{()]*)\)\s*" + terminator, _strip_comments(text), flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    return out


def test_ctypes_call_sites_pass_as_many_arguments_as_the_c_prototypes_take():
    """ctypes does not check argument counts: a call site that lags behind a changed signature would pass garbage.  Every
    `lib.dgb_*(...)` call in the tree is compared with the header, and the header with the definitions in `csrc/`."""
    import ast

    protos = _param_counts(open(os.path.join(ROOT, "include", "dgraph_b200.h")).read(), ";")
    assert len(protos) >= 50
    defs = {}
    for f in glob.glob(os.path.join(ROOT, "csrc", "*.cu")):
        defs.update(_param_counts(open(f).read(), r"\{"))
    assert {k: (v, defs.get(k)) for k, v in protos.items() if defs.get(k) != v} == {}
    bad = []
    for f in glob.glob(os.path.join(ROOT, "**", "*.py"), recursive=True):
        rel = os.path.relpath(f, ROOT)
        if rel.startswith(("baseline/_ref", "gpurun_out", "build")):
            continue
        for node in ast.walk(ast.parse(open(f).read())):
            if (isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and node.func.attr.startswith("dgb_")
                    and not any(isinstance(a, ast.Starred) for a in node.args) and not node.keywords):
                want = protos.get(node.func.attr)
                if want is None or want != len(node.args):
                    bad.append((rel, node.lineno, node.func.attr, len(node.args), want))
    assert not bad, bad


def test_ctypes_call_sites_pass_the_kinds_of_arguments_the_prototypes_take():
    """int vs int64_t vs float vs pointer, for every argument written as a recognisable wrapper (`N.ptr`, `N.i32`, `N.i64`, `N.f32`,
    `N.stream`, `ctypes.c_*`): a float handed to an integer slot (or the reverse) travels in the wrong register class."""
    import ast

    def kind(param):
        if "*" in param:
            return "ptr"
        base = re.sub(r"\b[A-Za-z_][A-Za-z0-9_]*$", "", param.strip()).replace("const", "").strip()
        return {"float": "f32", "double": "f64", "int": "i32", "unsigned": "i32", "unsigned int": "i32", "uint32_t": "i32",
                "int32_t": "i32", "int64_t": "i64", "long long": "i64", "size_t": "i64", "uint64_t": "i64"}.get(base, "?" + base)

    protos = {}
    hdr = _strip_comments(open(os.path.join(ROOT, "include", "dgraph_b200.h")).read())
    for m in re.finditer(r"\b(dgb_[a-z0-9_]+)\s*\(([^;{()]*)\)\s*;", hdr, flags=re.S):
        args = m.group(2).strip()
        protos[m.group(1)] = [] if args in ("", "void") else [kind(a) for a in args.split(",")]
    assert not {k for v in protos.values() for k in v if k.startswith("?")}
    wrappers = {"ptr": "ptr", "stream": "ptr", "byref": "ptr", "c_void_p": "ptr", "c_char_p": "ptr", "i32": "i32", "c_int": "i32",
                "c_uint": "i32", "i64": "i64", "c_longlong": "i64", "c_int64": "i64", "c_uint64": "i64", "c_size_t": "i64",
                "f32": "f32", "c_float": "f32", "c_double": "f64"}
    bad = []
    for f in glob.glob(os.path.join(ROOT, "**", "*.py"), recursive=True):
        rel = os.path.relpath(f, ROOT)
        if rel.startswith(("baseline/_ref", "gpurun_out", "build")):
            continue
        for node in ast.walk(ast.parse(open(f).read())):
            if isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and node.func.attr in protos:
                for i, (a, want) in enumerate(zip(node.args, protos[node.func.attr])):
                    if isinstance(a, ast.Call):
                        fn = a.func.attr if isinstance(a.func, ast.Attribute) else getattr(a.func, "id", "")
                        got = wrappers.get(fn)
                        if got is not None and got != want:
                            bad.append((rel, node.lineno, node.func.attr, i, got, want))
    assert not bad, bad


def test_host_header_declares_every_host_export_with_the_definitions_parameter_count():
    lib = os.path.join(ROOT, "dgraph_b200", "lib", "libdgraph_b200_host.so")
    if not os.path.exists(lib):
        pytest.skip("host library not built")
    hdr = _strip_comments(open(os.path.join(ROOT, "include", "dgraph_b200_host.h")).read())
    decl = {m.group(1): (0 if m.group(2).strip() in ("", "void") else len(m.group(2).split(",")))
            for m in re.finditer(r"\b(dgbh?_[a-z0-9_]+)\s*\(([^;{()]*)\)\s*;", hdr, flags=re.S)}
    exports = {s for s in _exports(lib) if s.startswith(("dgb_", "dgbh_"))}
    assert exports == set(decl), (exports ^ set(decl))
    defs = {}
    for f in glob.glob(os.path.join(ROOT, "csrc", "host", "*.cpp")):
        for m in re.finditer(r"\b(dgbh?_[a-z0-9_]+)\s*\(([^;{()]*)\)\s*\{", _strip_comments(open(f).read()), flags=re.S):
            defs[m.group(1)] = 0 if m.group(2).strip() in ("", "void") else len(m.group(2).split(","))
    assert {k: (v, defs.get(k)) for k, v in decl.items() if defs.get(k) != v} == {}
