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


# ---- the reference's helper names (``experiments/GraphCast/preprocess.py:17-87``), on networkx graphs -------------------------
def create_graphcast_meshgraph(mesh_level: int = 6):
    """``(vertices [V, 3], edges [E, 2])`` of the bidirected multimesh (every refinement level merged)."""
    from dgraph_b200.data.graphcast_graph import multimesh

    v, _faces, e, _max_len = multimesh(mesh_level)
    return torch.from_numpy(v).float(), torch.from_numpy(e).long()


def graphcast_graph_to_nxgraph(mesh_graph):
    import networkx as nx

    G = nx.Graph()
    G.add_nodes_from(range(int(mesh_graph[0].shape[0])))
    G.add_edges_from(map(tuple, mesh_graph[1].tolist()))
    return G


def save_nx_meshgraph(nx_graph, filename):
    import pickle

    with open(filename, "wb") as f:
        pickle.dump(nx_graph, f)


def partition_graph(G, num_ranks: int):
    """Vertex placement ``[num_nodes]`` of a networkx graph (METIS in the reference; the native multilevel partitioner here)."""
    from dgraph_b200.data.partition import multilevel

    e = torch.tensor(list(G.edges()), dtype=torch.long).reshape(-1, 2)
    return multilevel(e, G.number_of_nodes(), num_ranks)


def node_renumbering(node_rank_placement, num_parts):
    """New -> old vertex list: old ids grouped by rank."""
    return torch.argsort(torch.as_tensor(node_rank_placement).long(), stable=True)


def edge_renumbering(edge_indices, renumbered_nodes):
    """``[E, 2]`` edges in old ids -> ``[2, E]`` in new ids (through the inverse of the new -> old list)."""
    new_to_old = torch.as_tensor(renumbered_nodes).long()
    old_to_new = torch.empty_like(new_to_old)
    old_to_new[new_to_old] = torch.arange(new_to_old.numel())
    e = torch.as_tensor(list(edge_indices) if not torch.is_tensor(edge_indices) else edge_indices).long().reshape(-1, 2)
    return torch.stack([old_to_new[e[:, 0]], old_to_new[e[:, 1]]])


def metis_partition_graph(nx_graph, num_ranks):
    """``(placement, new -> old vertex list, renumbered edges)`` of the mesh graph (the reference's function stops after the
    renumbering, ``preprocess.py:76-87``; grid placement and bipartite edges are handled by
    ``DistributedGraphCastGraphGenerator(placement="multilevel")``)."""
    placement = partition_graph(nx_graph, num_ranks)
    nodes = node_renumbering(placement, num_ranks)
    return placement, nodes, edge_renumbering(list(nx_graph.edges()), nodes)


def main(load_graph: str = "", mesh_level: int = 6, num_ranks: int = 0, world_size: int = 4, lat: int = 721, lon: int = 1440,
         out_dir: str = "graphcast_cache", placement: str = "multilevel"):
    """``load_graph`` / ``mesh_level`` / ``num_ranks`` are the reference's arguments (``preprocess.py:89``); the mesh is always
    rebuilt here (a fraction of a second), so a pickled graph is not needed and ``load_graph`` is ignored.  ``num_ranks`` is an alias of
    ``world_size``."""
    world_size = int(num_ranks) or world_size
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
