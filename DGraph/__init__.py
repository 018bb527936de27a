"""Import-compatible façade: ``import DGraph`` resolves to the dgraph_b200 implementation.

Every public name of the reference package is re-exported from here so existing user code keeps working:
``DGraph.Communicator``, ``DGraph.distributed.{HaloExchange, DGraphMessagePassing, CommunicationPattern,
build_communication_pattern}``, ``DGraph.distributed.nccl.{NCCLBackendEngine, TIMINGS, NCCLGraphCommPlan, ...}``,
``DGraph.data.*``, ``DGraph.utils.TimingReport``.
"""
# The package attributes ``Communicator`` / ``CommunicatorBase`` are classes although submodules of the same names exist (as in
# the reference).  Importing a submodule for the first time rebinds the package attribute to the module, so both submodules are
# loaded here, before the classes are bound: a later ``import DGraph.CommunicatorBase`` then changes nothing.
from DGraph.Communicator import Communicator
from DGraph.CommunicatorBase import CommunicatorBase
from dgraph_b200.__version__ import __version__

__all__ = ["Communicator", "CommunicatorBase", "__version__"]
