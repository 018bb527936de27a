from dgraph_b200.communicator_base import CommunicatorBase  # noqa: F401
