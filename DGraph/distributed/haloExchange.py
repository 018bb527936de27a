from dgraph_b200.parallel.halo import HaloExchange, HaloExchangeImpl, DGraphMessagePassing  # noqa: F401
