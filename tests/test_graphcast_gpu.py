"""GraphCast on the GPU kernel path (bf16 gather-GEMM + segment sum) vs the fp32 PyTorch path of the same model."""
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def test_graphcast_bf16_kernel_path_matches_fp32_reference():
    from dgraph_b200.communicator import Communicator
    from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator, SyntheticWeatherDataset
    from dgraph_b200.models.graphcast import Config, DGraphCast
    from dgraph_b200.ops import _native
    from dgraph_b200.parallel.local_engine import LocalEngine

    torch.manual_seed(0)
    cfg = Config()
    cfg.model.mesh_level, cfg.data.latlon_res = 3, (46, 90)
    comm = Communicator._from_engine(LocalEngine(), backend="single")
    gen = DistributedGraphCastGraphGenerator(cfg.data.latlon_res, cfg.model.mesh_level)
    sg32 = gen.shard(0, 1).to("cuda", torch.float32)
    sg16 = gen.shard(0, 1).to("cuda", torch.bfloat16)
    model = DGraphCast(cfg, comm).cuda()
    x = SyntheticWeatherDataset(73, cfg.data.latlon_res)[0]["invar"].cuda()
    l0 = _native.launches()
    y16 = model(x.bfloat16(), sg16)
    assert _native.launches() > l0, "native kernels must run on the bf16 CUDA path"
    y32 = model(x, sg32)
    err = (y16.float() - y32).norm() / y32.norm()
    assert err < 3e-2, float(err)
    (y16.float() ** 2).mean().backward()
    g16 = torch.cat([p.grad.reshape(-1) for p in model.parameters()]).clone()
    model.zero_grad()
    (y32 ** 2).mean().backward()
    g32 = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    assert (g16 - g32).norm() / g32.norm() < 8e-2
