"""dgraph_b200 -- Blackwell-native distributed full-graph GNN training (DGraph capabilities, B200-first design)."""
from dgraph_b200.__version__ import __version__
from dgraph_b200.communicator_base import CommunicatorBase
from dgraph_b200.communicator import Communicator

__all__ = ["Communicator", "CommunicatorBase", "__version__"]
