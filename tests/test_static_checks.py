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
