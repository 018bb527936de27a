"""``DGraph.Communicator`` is both a submodule and (re-exported from the package) a class in the reference; importing the
submodule rebinds ``DGraph.Communicator`` to the module.  This module therefore forwards calls and attribute access to the
class, so ``from DGraph import Communicator; Communicator.init_process_group(...)`` works in either import order."""
import sys
import types

from dgraph_b200.communicator import Communicator, SUPPORTED_BACKENDS  # noqa: F401

# The reference's back-end tests write ``import DGraph.Communicator as Comm; Comm.Communicator.init_process_group(...)``
# (``tests/test_mpi_backend.py:2,16``, ``test_nccl_backend.py:17``).  On Python >= 3.7 that import binds ``Comm`` to the package
# attribute, i.e. to the class, so the spelling needs the class to be reachable from itself.
Communicator.Communicator = Communicator


class _ClassLikeModule(types.ModuleType):
    def __call__(self, *a, **k):
        return Communicator(*a, **k)

    def __getattr__(self, name):
        return getattr(Communicator, name)

    def __setattr__(self, name, value):
        if name.startswith("__") or name in ("Communicator", "SUPPORTED_BACKENDS"):
            super().__setattr__(name, value)
        else:
            setattr(Communicator, name, value)


sys.modules[__name__].__class__ = _ClassLikeModule
