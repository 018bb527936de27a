"""GraphCast: graph construction sizes, model parameter count / state_dict keys, distributed == single-process."""
import pytest
import torch

from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator, multimesh, SyntheticWeatherDataset
from dgraph_b200.models.graphcast import Config, DGraphCast
from dgraph_b200.parallel.emulator import run_world


def test_multimesh_sizes_match_reference():
    v, f, e, _ = multimesh(2)
    assert v.shape[0] == 10 * 4 ** 2 + 2 and f.shape[0] == 20 * 4 ** 2
    assert e.shape[0] == 4 * 30 * (1 + 4 + 16)  # all levels, both directions, duplicated like the reference
    v6, _, e6, _ = multimesh(6)
    assert v6.shape[0] == 40962 and e6.shape[0] == 655320  # experiments/GraphCast/tests/test_single_graph_data.py:21-33


def test_model_parameter_count_and_keys():
    m = DGraphCast(Config(), comm=None)
    assert sum(p.numel() for p in m.parameters()) == 1025627
    assert len(list(m.parameters())) == 114
    sd = m.state_dict()
    assert tuple(sd["processor.edge_processors.0.mesh_mlp._model.0.weight"].shape) == (128, 384)
    assert tuple(sd["processor.node_processors.3.mesh_mlp._model.0.weight"].shape) == (128, 256)
    assert "embedder.grid_feature_embedder._model.3.weight" in sd and "final_prediction._model.2.bias" in sd


def _cfg():
    c = Config()
    c.model.hidden_dim, c.model.processor_layers, c.model.mesh_level = 16, 2, 2
    c.model.input_grid_dim = c.model.output_grid_dim = 5
    c.data.latlon_res = (13, 24)
    return c


@pytest.mark.parametrize("W,placement", [(2, "bands"), (3, "bands"), (3, "multilevel")])
def test_distributed_graphcast_matches_single_process(W, placement):
    torch.manual_seed(0)
    cfg = _cfg()
    gen = DistributedGraphCastGraphGenerator(cfg.data.latlon_res, cfg.model.mesh_level, placement=placement)
    ref = DGraphCast(cfg, comm=None).double()
    state = {k: v.clone() for k, v in ref.state_dict().items()}
    x_full = SyntheticWeatherDataset(5, cfg.data.latlon_res, dtype=torch.float64)[0]["invar"]

    def job(comm):
        r, w = comm.get_rank(), comm.get_world_size()
        sg = gen.shard(r, w).to("cpu", torch.float64)
        model = DGraphCast(cfg, comm).double()
        model.load_state_dict(state)
        x = x_full[sg.grid_rows_global].clone().requires_grad_(True)
        y = model(x, sg)
        (y ** 2).sum().backward()
        g = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
        comm.all_reduce(g)
        return sg.grid_rows_global, y.detach(), x.grad, g

    rows1, y1, gx1, g1 = run_world(1, job)[0]
    for rows, y, gx, g in run_world(W, job):
        assert torch.allclose(y, y1[rows], atol=1e-8)
        assert torch.allclose(gx, gx1[rows], atol=1e-8)
        assert torch.allclose(g, g1, atol=1e-7)
