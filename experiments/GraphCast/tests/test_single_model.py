"""One forward of the single-process model (reference ``tests/test_single_model.py``): output shape and finiteness."""
import torch


def test_forward_shape(graph_generator, graphcast_config):
    from dgraph_b200.models.graphcast import DGraphCast

    cfg = graphcast_config
    sg = graph_generator.shard(0, 1)
    torch.manual_seed(0)
    model = DGraphCast(cfg, comm=None)
    x = torch.randn(sg.num_local_grid, cfg.model.input_grid_dim)
    with torch.no_grad():
        y = model(x, sg)
    assert y.shape == (sg.num_local_grid, cfg.model.output_grid_dim)
    assert bool(torch.isfinite(y).all())
