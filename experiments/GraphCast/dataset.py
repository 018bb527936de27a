"""Reference file name (``experiments/GraphCast/dataset.py:24-226``).

``SyntheticWeatherDataset`` here takes the reference's constructor arguments (``channels`` as a list, ``num_samples_per_year``,
``num_steps``, optional ``mesh_vertex_placement``, ``device``, ``grid_size``, ``mesh_level``, ``rank`` / ``world_size`` /
``ranks_per_graph``, free keyword arguments) and offers ``get_static_graph()``; the fields and the graph come from
``dgraph_b200.data.graphcast_graph``."""
import os
import sys
from typing import Any, List, Optional, Sequence, Tuple, Union

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

import torch  # noqa: E402

from dgraph_b200.data.graphcast_graph import (DistributedGraphCastGraph, DistributedGraphCastGraphGenerator,  # noqa: F401,E402
                                              SyntheticWeatherDataset as _Fields)

__all__ = ["SyntheticWeatherDataset", "DistributedGraphCastGraphGenerator", "DistributedGraphCastGraph"]


class SyntheticWeatherDataset(torch.utils.data.Dataset):
    def __init__(self, channels: Union[int, Sequence[int]], num_samples_per_year: int, num_steps: int = 1,
                 mesh_vertex_placement: Optional[torch.Tensor] = None, device: Union[str, torch.device] = "cpu",
                 grid_size: Tuple[int, int] = (721, 1440), base_temp: float = 15, amplitude: float = 10, noise_level: float = 2,
                 mesh_level: int = 6, rank: int = 0, world_size: int = 1, ranks_per_graph: int = 1, **kwargs: Any):
        self.num_channels = channels if isinstance(channels, int) else len(channels)
        self.num_days, self.num_steps, self.device = int(num_samples_per_year), int(num_steps), torch.device(device)
        self.grid_size, self.mesh_level = tuple(grid_size), mesh_level
        self.rank, self.world_size, self.ranks_per_graph = rank, world_size, max(int(ranks_per_graph), 1)
        self.extra_args = kwargs
        self.base_temp, self.amplitude, self.noise_level = base_temp, amplitude, noise_level
        # ranks of one replica share a graph.  `mesh_vertex_placement` (the per-vertex rank tensor of a partitioner run, a file in
        # the reference) is used as given -- the grid follows its mesh neighbours; without it the graph is cut in latitude bands
        self._gen = DistributedGraphCastGraphGenerator(self.grid_size, mesh_level)
        g_rank, g_world = rank % self.ranks_per_graph, self.ranks_per_graph if world_size > 1 else 1
        if mesh_vertex_placement is not None and g_world > 1:
            graph = self._gen.get_graphcast_graph(torch.as_tensor(mesh_vertex_placement), rank=g_rank, world_size=g_world)
        else:
            graph = self._gen.shard(g_rank, g_world)
        self._graph = graph.to(self.device, torch.float32)
        self._fields = _Fields(self.num_channels, self.grid_size, num_samples=self.num_days,
                               grid_rows=self._graph.grid_rows_global.cpu())

    def generate_data(self, num_days: int, num_channels: int, grid_size: Tuple[int, int], base_temp: float, amplitude: float,
                      noise_level: float):
        """Dense ``[days, channels, lat, lon]`` field of the reference's generator (``dataset.py:100-171``).  ``__getitem__`` does
        not go through it: samples are produced row by row for this rank's grid points only."""
        return _Fields.generate_data(num_days, num_channels, grid_size, base_temp, amplitude, noise_level)

    def get_static_graph(self) -> DistributedGraphCastGraph:
        return self._graph

    def __len__(self) -> int:
        return self.num_days

    def __getitem__(self, idx: int):
        item = self._fields[idx]
        return {"invar": item["invar"].to(self.device), "outvar": item["outvar"].to(self.device)}
