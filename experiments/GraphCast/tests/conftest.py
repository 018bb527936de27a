"""Session fixtures of the GraphCast integration tests (parity: reference ``experiments/GraphCast/tests/conftest.py``).

``python -m pytest experiments/GraphCast/tests`` builds the graph once per session.  The full 0.25 degree graph (721 x 1440 grid,
mesh level 6) takes about a minute on the CPU; set ``DGRAPH_GRAPHCAST_FULL=0`` for a coarse graph (mesh level 3, 46 x 90 grid).
``DGRAPH_GRAPHCAST_DISTRIBUTED_TESTS`` is accepted for compatibility (distributed cases live in ``tests/test_graphcast.py``)."""
import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import _common  # noqa: F401,E402

FULL = os.environ.get("DGRAPH_GRAPHCAST_FULL", "1") == "1"


@pytest.fixture(scope="session")
def graphcast_config():
    from dgraph_b200.models.graphcast import Config

    cfg = Config()
    if not FULL:
        cfg.model.mesh_level, cfg.data.latlon_res = 3, (46, 90)
    return cfg


@pytest.fixture(scope="session")
def graph_generator(graphcast_config):
    from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator

    gen = DistributedGraphCastGraphGenerator(graphcast_config.data.latlon_res, graphcast_config.model.mesh_level)
    gen.global_graph()
    return gen
