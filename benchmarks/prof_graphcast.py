"""One-GPU GraphCast train-step driver for ncu / timing (full 0.25 degree graph unless LAT/LON/LEVEL are set)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgraph_b200.communicator import Communicator  # noqa: E402
from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator, SyntheticWeatherDataset  # noqa: E402
from dgraph_b200.models.graphcast import Config, DGraphCast  # noqa: E402
from dgraph_b200.optim import FusedAdam  # noqa: E402
from dgraph_b200.parallel.local_engine import LocalEngine  # noqa: E402

cfg = Config()
cfg.data.latlon_res = (int(os.environ.get("LAT", 721)), int(os.environ.get("LON", 1440)))
cfg.model.mesh_level = int(os.environ.get("LEVEL", 6))
steps = int(os.environ.get("STEPS", 3))
comm = Communicator._from_engine(LocalEngine(), backend="single")
sg = DistributedGraphCastGraphGenerator(cfg.data.latlon_res, cfg.model.mesh_level).shard(0, 1).to("cuda", torch.bfloat16)
d = SyntheticWeatherDataset(73, cfg.data.latlon_res)[0]
x, y = d["invar"].cuda().bfloat16(), d["outvar"].cuda().bfloat16()
model = DGraphCast(cfg, comm).cuda()
opt = FusedAdam(model.parameters(), lr=1e-3)


def step():
    opt.zero_grad()
    loss = torch.nn.functional.mse_loss(model(x, sg).float(), y.float())
    loss.backward()
    opt.step()
    return loss


for _ in range(2):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(steps):
    l = step()
torch.cuda.synchronize()
print(f"graphcast step {1e3 * (time.perf_counter() - t0) / steps:.2f} ms loss {float(l.detach()):.5f}")
