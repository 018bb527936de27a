from dgraph_b200.__version__ import __version__  # noqa: F401
