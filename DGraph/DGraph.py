"""Empty in the reference as well (``DGraph/DGraph.py``); kept so that the module path resolves."""
