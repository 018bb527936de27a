"""Build and store the GraphCast graphs + vertex placements (parity: reference ``experiments/GraphCast/preprocess.py:17-101``
and ``data_utils/preprocess.py``; METIS is replaced by latitude bands aligned between grid and mesh).

``python experiments/GraphCast/preprocess.py --world_size 4 --out_dir graphcast_cache``"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

import torch  # noqa: E402

from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator  # noqa: E402
from dgraph_b200.utils.cli import run_cli  # noqa: E402


def main(world_size: int = 4, mesh_level: int = 6, lat: int = 721, lon: int = 1440, out_dir: str = "graphcast_cache", placement: str = "multilevel"):
    gen = DistributedGraphCastGraphGenerator((lat, lon), mesh_level, placement=placement)
    g = gen.global_graph()
    grid_pl, mesh_pl = gen.placements(world_size)
    os.makedirs(out_dir, exist_ok=True)
    torch.save(mesh_pl, os.path.join(out_dir, f"mesh_vertex_rank_placement_{world_size}.pt"))
    torch.save(grid_pl, os.path.join(out_dir, f"grid_vertex_rank_placement_{world_size}.pt"))
    print(f"grid {g.num_grid} mesh {g.num_mesh} | edges mesh {g.mesh_edges.shape[0]} g2m {g.grid2mesh_edges.shape[0]} "
          f"m2g {g.mesh2grid_edges.shape[0]} | placements for {world_size} ranks -> {out_dir}")
    for r in range(world_size):
        s = gen.shard(r, world_size)
        print(f"  rank {r}: grid {s.num_local_grid} mesh {s.num_local_mesh} halo g2m {s.patterns.grid2mesh.num_halo_vertices} "
              f"mesh {s.patterns.mesh.num_halo_vertices} m2g {s.patterns.mesh2grid.num_halo_vertices}")


if __name__ == "__main__":
    run_cli(main)
