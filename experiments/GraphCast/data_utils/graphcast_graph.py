"""Reference file name (``data_utils/graphcast_graph.py:34-437``)."""
from dgraph_b200.data.graphcast_graph import (DistributedGraphCastGraph, DistributedGraphCastGraphGenerator,  # noqa: F401
                                              GraphCastCommPatterns, GraphCastGraph, build_graphcast_comm_patterns)


from dataclasses import dataclass  # noqa: E402

from torch import Tensor  # noqa: E402


@dataclass
class GraphCastTopology:
    """Index-only description of the three GraphCast graphs on one rank (reference ``graphcast_graph.py:42-56``): the input of
    ``build_graphcast_comm_patterns`` there.  Here the generator builds the patterns itself
    (``DistributedGraphCastGraphGenerator.shard``); ``from_generator`` fills this container from it for code that wants the raw
    placements and global edge lists."""
    rank: int
    world_size: int
    ranks_per_graph: int
    mesh_rank_placement: Tensor
    grid_rank_placement: Tensor
    mesh_graph_src_indices: Tensor
    mesh_graph_dst_indices: Tensor
    mesh2grid_graph_src_indices: Tensor
    mesh2grid_graph_dst_indices: Tensor
    grid2mesh_graph_src_indices: Tensor
    grid2mesh_graph_dst_indices: Tensor

    @staticmethod
    def from_generator(generator: "DistributedGraphCastGraphGenerator", rank: int, world_size: int,
                       ranks_per_graph: int = -1) -> "GraphCastTopology":
        g = generator.global_graph()
        grid_pl, mesh_pl = generator.placements(world_size)
        return GraphCastTopology(
            rank=rank, world_size=world_size, ranks_per_graph=world_size if ranks_per_graph < 0 else ranks_per_graph,
            mesh_rank_placement=mesh_pl, grid_rank_placement=grid_pl,
            mesh_graph_src_indices=g.mesh_edges[:, 0], mesh_graph_dst_indices=g.mesh_edges[:, 1],
            mesh2grid_graph_src_indices=g.mesh2grid_edges[:, 0], mesh2grid_graph_dst_indices=g.mesh2grid_edges[:, 1],
            grid2mesh_graph_src_indices=g.grid2mesh_edges[:, 0], grid2mesh_graph_dst_indices=g.grid2mesh_edges[:, 1])
