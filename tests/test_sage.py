import pytest
import torch

from dgraph_b200.models.sage import CommAwareSAGE
from dgraph_b200.parallel import HaloExchange, build_communication_pattern
from dgraph_b200.parallel.emulator import run_world
from tests._graphs import random_graph, partition


@pytest.mark.parametrize("W", [1, 3])
def test_sage_distributed_matches_dense_oracle(W):
    torch.manual_seed(0)
    V, F = 50, 6
    e = random_graph(V, 5, seed=2, symmetric=False)
    xg = torch.randn(V, F, dtype=torch.float64)
    ref = CommAwareSAGE(F, 8, 3, None, None).double()
    state = ref.state_dict()
    # dense oracle
    A = torch.zeros(V, V, dtype=torch.float64)
    A[e[:, 0], e[:, 1]] = 1
    A = A / A.sum(1, keepdim=True).clamp(min=1)
    h = xg
    for conv in ref.convs:
        h = torch.relu(torch.cat([h, A @ h], 1) @ conv.lin.weight.t() + conv.lin.bias)
    want = h @ ref.fc.weight.t() + ref.fc.bias
    part = partition(V, W, "random")

    def job(comm):
        pat = build_communication_pattern(e, part, comm.get_rank(), W)
        m = CommAwareSAGE(F, 8, 3, HaloExchange(comm), comm).double()
        m.load_state_dict(state)
        return pat.local_vertices_global, m(xg[pat.local_vertices_global], pat)

    for gids, out in run_world(W, job):
        assert torch.allclose(out, want[gids], atol=1e-9)
