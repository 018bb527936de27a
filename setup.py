"""``pip install -e .`` / ``python setup.py build_ext --inplace``: compiles csrc/*.cu for sm_100a with the in-tree nvcc driver
(``dgraph_b200/build.py``) -- the counterpart of the reference's scikit-build-core + CMake flow (``pyproject.toml:1-66``,
``CMakeLists.txt:1-192``).  ``CMakeLists.txt`` next to this file builds the same library for CMake users."""
import importlib.util
import os

from setuptools import setup
from setuptools.command.build_ext import build_ext
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))


def _native_build():
    spec = importlib.util.spec_from_file_location("_dgb_build", os.path.join(HERE, "dgraph_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    print("dgraph_b200: built", mod.build())
    try:  # host-only C++ (partitioner, shared-memory windows); the Python side has fallbacks without it
        print("dgraph_b200: built", mod.build_host())
    except Exception as e:  # noqa: BLE001
        print("dgraph_b200: host library not built:", e)


class BuildPy(build_py):
    def run(self):
        if os.environ.get("DGRAPH_B200_SKIP_NATIVE", "0") != "1":
            _native_build()
        super().run()


class BuildExt(build_ext):
    def run(self):
        _native_build()


setup(cmdclass={"build_py": BuildPy, "build_ext": BuildExt})
