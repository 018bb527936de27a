"""Import-compatible façade: ``import DGraph`` resolves to the dgraph_b200 implementation.

Every public name of the reference package is re-exported from here so existing user code keeps working:
``DGraph.Communicator``, ``DGraph.distributed.{HaloExchange, DGraphMessagePassing, CommunicationPattern,
build_communication_pattern}``, ``DGraph.distributed.nccl.{NCCLBackendEngine, TIMINGS, NCCLGraphCommPlan, ...}``,
``DGraph.data.*``, ``DGraph.utils.TimingReport``.
"""
from dgraph_b200.communicator import Communicator
from dgraph_b200.communicator_base import CommunicatorBase
from dgraph_b200.__version__ import __version__

__all__ = ["Communicator", "CommunicatorBase", "__version__"]
