"""Graph sizes of the single-process dataset (reference ``tests/test_single_graph_data.py:10-33``)."""
from conftest import FULL


def test_graph_sizes(graph_generator, graphcast_config):
    g = graph_generator.global_graph()
    n_lat, n_lon = graphcast_config.data.latlon_res
    assert g.num_grid == n_lat * n_lon
    assert g.mesh_edges.shape[1] == 2 and g.mesh_edge_features.shape == (g.mesh_edges.shape[0], 4)
    assert g.mesh_node_features.shape == (g.num_mesh, 3)
    assert g.mesh2grid_edges.shape[0] == 3 * g.num_grid  # three corners of the containing face per grid point
    if FULL:  # the numbers the reference asserts for 0.25 degrees / mesh level 6
        assert g.num_grid == 1_038_240 and g.num_mesh == 40_962
        assert g.mesh_edges.shape[0] == 655_320
        assert g.mesh2grid_edges.shape[0] == 3_114_720
        # grid2mesh: 1,618,824 in the reference (1,618,746 in the paper, data_utils/utils.py:160).  With the reference's mesh
        # orientation (a face, not an edge, perpendicular to the polar axis) the count is reproduced; the neighbour-radius test is
        # floating-point sensitive at the boundary, so a handful of edges of slack is allowed
        assert abs(g.grid2mesh_edges.shape[0] - 1_618_824) <= 64, g.grid2mesh_edges.shape[0]


def test_single_rank_shard_is_the_whole_graph(graph_generator):
    g = graph_generator.global_graph()
    sg = graph_generator.shard(0, 1)
    assert sg.num_local_grid == g.num_grid and sg.num_local_mesh == g.num_mesh
    assert sg.mesh_graph_src_indices.numel() == g.mesh_edges.shape[0]
    assert sg.patterns.mesh.num_halo_vertices == 0
