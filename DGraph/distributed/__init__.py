from dgraph_b200.parallel.halo import HaloExchange, HaloExchangeImpl, DGraphMessagePassing
from dgraph_b200.parallel.comm_pattern import CommunicationPattern, build_communication_pattern

__all__ = ["HaloExchange", "HaloExchangeImpl", "DGraphMessagePassing", "CommunicationPattern", "build_communication_pattern"]
