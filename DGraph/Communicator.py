from dgraph_b200.communicator import Communicator, SUPPORTED_BACKENDS  # noqa: F401
