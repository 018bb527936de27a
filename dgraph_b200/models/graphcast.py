"""GraphCast encoder-processor-decoder mesh GNN.

Parity: reference ``experiments/GraphCast/{layers,model,graphcast_config}.py`` -- ``MeshGraphMLP`` (Linear, SiLU, [Linear,
SiLU]*, Linear, LayerNorm), ``MeshEdgeBlock`` (``MLP(cat[src, dst, e]) + e``), ``MeshNodeBlock`` (``MLP(cat[n, sum(e)]) + n``),
``GraphCastEmbedder / Encoder / Processor / Decoder``, ``DGraphCast`` with identical module names, hence identical
``state_dict`` keys (e.g. ``processor.edge_processors.0.mesh_mlp._model.0.weight`` of shape ``(128, 384)``); 114 tensors /
1,025,627 parameters at the default config.

B200 execution (CUDA bf16): the first Linear of every edge / node block is a tcgen05 *gather-GEMM* over the un-gathered,
un-concatenated sources (``ops.gather_linear``); the scatter-sum of a node block is a contiguous segment sum because edges
are stored sorted by their aggregating vertex; the only communication of an edge block is ONE halo pull of the neighbour
side's node features (``HaloExchange.augmented``).  On CPU / fp32 the same modules run plain PyTorch.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from dgraph_b200.ops import mlp_fused as _mlp_fused
from dgraph_b200.ops.gather_linear import gather_linear
from dgraph_b200.ops.gemm import linear
from dgraph_b200.ops.norm import layer_norm_residual, silu
from dgraph_b200.ops.segment import segment_sum
from dgraph_b200.parallel.halo import HaloExchange
from dgraph_b200.parallel.kernel_plan import GraphKernelPlan


# ------------------------------------------------------------------------------------------------ config
@dataclass
class TrainingConfig:
    lr: float = 1e-3
    lr_step3: float = 3e-7
    num_iters_step1: int = 1000
    num_iters_step2: int = 299000
    num_iters_step3: int = 11000
    step_change_freq: int = 1000
    save_freq: int = 1
    grad_clip_norm: float = 32.0
    val_freq: int = 5


@dataclass
class DataConfig:
    latlon_res: Tuple[int, int] = (721, 1440)
    num_samples_per_year_train: int = 4
    num_channels_climate: int = 73
    num_channels_static: int = 5
    num_history: int = 0
    use_cos_zenith: bool = True
    dt: float = 6.0
    start_year: int = 1980
    use_time_of_year_index: bool = True
    stride: int = 1


@dataclass
class ModelConfig:
    processor_layers: int = 4
    hidden_dim: int = 128
    mesh_level: int = 6
    multimesh: bool = True
    input_grid_dim: int = 73
    input_mesh_dim: int = 3
    input_edge_dim: int = 4
    output_grid_dim: int = 73


@dataclass
class Config:
    training: TrainingConfig = field(default_factory=TrainingConfig)
    data: DataConfig = field(default_factory=DataConfig)
    model: ModelConfig = field(default_factory=ModelConfig)


# ------------------------------------------------------------------------------------------------ layers
def _pad8(x: torch.Tensor) -> torch.Tensor:
    """bf16 kernels want rows that are multiples of 16 bytes: zero-pad narrow inputs (3-, 4-, 73-wide) once."""
    k = x.shape[1]
    if x.is_cuda and x.dtype == torch.bfloat16 and k % 8:
        return F.pad(x, (0, 8 - k % 8))
    return x


class MeshGraphMLP(nn.Module):
    def __init__(self, input_dim: int, output_dim: int, hidden_dim: int = 512, hidden_layers: int = 1,
                 activation_fn: nn.Module = None, norm_type: str = "LayerNorm"):
        super().__init__()
        self.input_dim, self.output_dim = input_dim, output_dim
        act = activation_fn if activation_fn is not None else nn.SiLU()
        layers = [nn.Linear(input_dim, hidden_dim), act]
        for _ in range(hidden_layers - 1):
            layers += [nn.Linear(hidden_dim, hidden_dim), act]
        layers.append(nn.Linear(hidden_dim, output_dim))
        layers.append(getattr(nn, norm_type)(output_dim))
        self._model = nn.Sequential(*layers)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.forward_sources([x], [None])

    def forward_sources(self, srcs, idxs, residual=None, scatter_plans=None) -> torch.Tensor:
        """MLP applied to ``cat([srcs[k][idxs[k]] for k], 1)`` -- the gather and the concat happen inside the first GEMM;
        ``residual`` is added inside the fused LayerNorm kernel."""
        mods = list(self._model)
        first = mods[0]
        fused_silu = isinstance(mods[1], nn.SiLU)
        widths = [s.shape[1] for s in srcs]
        srcs = [_pad8(s) for s in srcs]
        w = first.weight
        if any(s.shape[1] != k for s, k in zip(srcs, widths)):  # inputs were padded: pad the matching weight columns
            parts, o = [], 0
            for s, k in zip(srcs, widths):
                parts.append(F.pad(w[:, o:o + k], (0, s.shape[1] - k)))
                o += k
            w = torch.cat(parts, dim=1)
        if (_mlp_fused.ENABLED and len(mods) == 4 and fused_silu and isinstance(mods[2], nn.Linear)
                and isinstance(mods[3], nn.LayerNorm) and _mlp_fused.supported(srcs, w, mods[2].weight, residual)):
            # default (DGRAPH_B200_FUSED_MLP=0 turns it off): the whole block in ONE kernel, h never leaves the SM (csrc/mlp_fused_tcgen05.cu)
            return _mlp_fused.mlp2_ln_fused(srcs, idxs, w, first.bias, mods[2].weight, mods[2].bias, mods[3].weight, mods[3].bias,
                                            mods[3].eps, residual, scatter_plans=scatter_plans)
        h = gather_linear(srcs, idxs, w, first.bias, act="silu" if fused_silu else None, scatter_plans=scatter_plans)
        i = 2 if fused_silu else 1
        used_res = False
        while i < len(mods):
            m = mods[i]
            if isinstance(m, nn.Linear):
                h = linear(h, m.weight.to(h.dtype), m.bias.to(h.dtype))
            elif isinstance(m, nn.LayerNorm):
                last = i == len(mods) - 1
                h = layer_norm_residual(h, m.weight, m.bias, m.eps, residual if last else None)
                used_res = used_res or last
            elif isinstance(m, nn.SiLU):
                h = silu(h)
            else:
                h = m(h)
            i += 1
        if residual is not None and not used_res:
            h = h + residual
        return h


def _edge_plan(static_graph, name: str, central_idx: torch.Tensor, n_central: int) -> GraphKernelPlan:
    """Kernel plan of the (central vertex <- edge id) incidence: node-block aggregation = segment sum over edge rows."""
    key = f"agg:{name}"
    kp = static_graph._plans.get(key)
    if kp is None:
        E = int(central_idx.shape[0])
        inc = torch.stack([central_idx.long(), torch.arange(E, device=central_idx.device)], dim=1)
        kp = GraphKernelPlan.from_edges(inc, n_central, E, device=central_idx.device, num_nbr_local=E)
        static_graph._plans[key] = kp
    return kp


class MeshNodeBlock(nn.Module):
    def __init__(self, input_node_dim: int, input_edge_dim: int, output_node_dim: int, comm, hidden_dim: int = 512,
                 num_hidden_layers: int = 1, aggregation_type: str = "sum"):
        super().__init__()
        assert aggregation_type in ["sum"], "Only sum aggregation is supported for now."
        self.aggregation_type, self.comm = aggregation_type, comm
        self.mesh_mlp = MeshGraphMLP(input_node_dim + input_edge_dim, output_node_dim, hidden_dim, num_hidden_layers)

    def forward(self, node_features, edge_features, src_indices, rank_mapping=None, *, agg_plan: Optional[GraphKernelPlan] = None):
        n = node_features.shape[0]
        if agg_plan is not None:
            agg = segment_sum(edge_features, agg_plan)
        elif rank_mapping is not None:
            agg = self.comm.scatter(edge_features, src_indices, rank_mapping, n)
        else:
            agg = torch.zeros((n, edge_features.shape[1]), dtype=edge_features.dtype, device=edge_features.device)
            agg = agg.index_add(0, src_indices.long(), edge_features)
        return self.mesh_mlp.forward_sources([node_features, agg], [None, None], residual=node_features)


class MeshEdgeBlock(nn.Module):
    def __init__(self, input_src_node_dim: int, input_dst_node_dim: int, input_edge_dim: int, output_edge_dim: int, comm,
                 hidden_dim: int = 512, num_hidden_layers: int = 1, aggregation_type: str = "sum"):
        super().__init__()
        assert aggregation_type in ["sum"], "Only sum aggregation is supported for now."
        self.aggregation_type, self.comm = aggregation_type, comm
        self.mesh_mlp = MeshGraphMLP(input_src_node_dim + input_dst_node_dim + input_edge_dim, output_edge_dim, hidden_dim,
                                     num_hidden_layers)

    def forward(self, src_node_features, dst_node_features, edge_features, src_indices, dst_indices, src_rank_mapping=None,
                dst_rank_mapping=None, *, scatter_plans=None):
        """``src_node_features`` / ``dst_node_features`` are the matrices the indices refer to (already ``[local | halo]``
        for the side that crosses ranks).  With rank mappings the reference's gather API is used instead."""
        if src_rank_mapping is not None or dst_rank_mapping is not None:
            s = self.comm.gather(src_node_features, src_indices, src_rank_mapping)
            d = self.comm.gather(dst_node_features, dst_indices, dst_rank_mapping)
            return self.mesh_mlp.forward_sources([s, d, edge_features], [None, None, None]) + edge_features
        return self.mesh_mlp.forward_sources([src_node_features, dst_node_features, edge_features],
                                             [src_indices, dst_indices, None], residual=edge_features,
                                             scatter_plans=scatter_plans)


class GraphCastEmbedder(nn.Module):
    def __init__(self, cfg: Config, *args, **kwargs):
        super().__init__()
        m = cfg.model
        self.grid_input_dim, self.mesh_input_dim, self.hidden_dim = m.input_grid_dim, m.input_mesh_dim, m.hidden_dim
        mk = lambda i: MeshGraphMLP(i, m.hidden_dim, m.hidden_dim, 1)  # noqa: E731
        self.grid_feature_embedder = mk(m.input_grid_dim)
        self.mesh_feature_embedder = mk(m.input_mesh_dim)
        self.grid2mesh_edge_embedder = mk(m.input_edge_dim)
        self.mesh2grid_edge_embedder = mk(m.input_edge_dim)
        self.mesh2mesh_edge_embedder = mk(m.input_edge_dim)

    def forward(self, grid_features, mesh_features, mesh2mesh_edge_features, grid2mesh_edge_features, mesh2grid_edge_features):
        return (self.grid_feature_embedder(grid_features), self.mesh_feature_embedder(mesh_features),
                self.mesh2mesh_edge_embedder(mesh2mesh_edge_features), self.grid2mesh_edge_embedder(grid2mesh_edge_features),
                self.mesh2grid_edge_embedder(mesh2grid_edge_features))


class GraphCastEncoder(nn.Module):
    def __init__(self, cfg: Config, comm, *args, **kwargs):
        super().__init__()
        h = cfg.model.hidden_dim
        self.edge_mlp = MeshEdgeBlock(h, h, h, h, comm, h)
        self.mesh_node_mlp = MeshNodeBlock(h, h, h, comm, h)
        self.grid_node_mlp = MeshGraphMLP(input_dim=h, output_dim=h)

    def forward(self, grid_node_features, mesh_node_features, grid2mesh_edge_features, grid2mesh_edge_indices_src,
                grid2mesh_edge_indices_dst, *, grid_aug=None, agg_plan=None, edge_scatter_plans=None):
        e = self.edge_mlp(grid_aug if grid_aug is not None else grid_node_features, mesh_node_features,
                          grid2mesh_edge_features, grid2mesh_edge_indices_src, grid2mesh_edge_indices_dst,
                          scatter_plans=edge_scatter_plans)
        n = self.mesh_node_mlp(mesh_node_features, e, grid2mesh_edge_indices_dst, agg_plan=agg_plan)
        mesh_node_features = mesh_node_features + n
        grid_node_features = grid_node_features + self.grid_node_mlp(grid_node_features)
        return grid_node_features, mesh_node_features


class GraphCastProcessor(nn.Module):
    def __init__(self, cfg: Config, comm, *args, **kwargs):
        super().__init__()
        h = cfg.model.hidden_dim
        self.edge_processors = nn.ModuleList([MeshEdgeBlock(h, h, h, h, comm, h) for _ in range(cfg.model.processor_layers)])
        self.node_processors = nn.ModuleList([MeshNodeBlock(h, h, h, comm, h) for _ in range(cfg.model.processor_layers)])

    def forward(self, embedded_mesh_features, embedded_mesh2mesh_edge_features, mesh2mesh_edge_indices_src,
                mesh2mesh_edge_indices_dst, *, exchange=None, agg_plan=None, edge_scatter_plans=None):
        e, n = embedded_mesh2mesh_edge_features, embedded_mesh_features
        for edge_layer, node_layer in zip(self.edge_processors, self.node_processors):
            n_aug = exchange(n) if exchange is not None else n  # dst endpoints may live on other ranks
            e = edge_layer(n, n_aug, e, mesh2mesh_edge_indices_src, mesh2mesh_edge_indices_dst,
                           scatter_plans=edge_scatter_plans)
            n = node_layer(n, e, mesh2mesh_edge_indices_src, agg_plan=agg_plan)
        return n, e


class GraphCastDecoder(nn.Module):
    def __init__(self, cfg: Config, comm, *args, **kwargs):
        super().__init__()
        h = cfg.model.hidden_dim
        self.comm = comm
        self.edge_mlp = MeshEdgeBlock(h, h, h, h, comm, h)
        self.node_mlp = MeshNodeBlock(input_node_dim=h, input_edge_dim=h, output_node_dim=h, hidden_dim=h, comm=comm,
                                      num_hidden_layers=1)

    def forward(self, mesh2grid_edge_features, grid_node_features, mesh_node_features, mesh2grid_edge_indices_src,
                mesh2grid_edge_indices_dst, *, mesh_aug=None, agg_plan=None, edge_scatter_plans=None):
        e = self.edge_mlp(mesh_aug if mesh_aug is not None else mesh_node_features, grid_node_features,
                          mesh2grid_edge_features, mesh2grid_edge_indices_src, mesh2grid_edge_indices_dst,
                          scatter_plans=edge_scatter_plans)
        n = self.node_mlp(grid_node_features, e, mesh2grid_edge_indices_dst, agg_plan=agg_plan)
        return grid_node_features + n


class DGraphCast(nn.Module):
    def __init__(self, cfg: Config, comm, *args, **kwargs):
        super().__init__()
        self.hidden_dim, self.output_grid_dim, self.comm = cfg.model.hidden_dim, cfg.model.output_grid_dim, comm
        self.embedder = GraphCastEmbedder(cfg=cfg, comm=comm)
        self.encoder = GraphCastEncoder(cfg=cfg, comm=comm)
        self.processor = GraphCastProcessor(cfg=cfg, comm=comm)
        self.decoder = GraphCastDecoder(cfg=cfg, comm=comm)
        self.final_prediction = MeshGraphMLP(input_dim=self.hidden_dim, output_dim=self.output_grid_dim)

    def forward(self, input_grid_features: torch.Tensor, static_graph) -> torch.Tensor:
        x_in = input_grid_features.squeeze(0) if input_grid_features.dim() == 3 else input_grid_features
        sg = static_graph
        pats = getattr(sg, "patterns", None)
        ex = HaloExchange(self.comm) if (pats is not None and self.comm is not None) else None
        halo = (lambda x, p: ex.augmented(x, p)) if ex is not None else (lambda x, p: x)
        dt = x_in.dtype
        grid, mesh, e_mm, e_g2m, e_m2g = self.embedder(x_in, sg.mesh_graph_node_features.to(dt),
                                                       sg.mesh_graph_edge_features.to(dt),
                                                       sg.grid2mesh_graph_edge_features.to(dt),
                                                       sg.mesh2grid_graph_edge_features.to(dt))
        n_mesh, n_grid = mesh.shape[0], grid.shape[0]
        use_plans = hasattr(sg, "_plans") and x_in.is_cuda and x_in.dtype == torch.bfloat16
        plan = lambda name, idx, n: _edge_plan(sg, name, idx, n) if use_plans else None  # noqa: E731
        grid_aug = halo(grid, pats.grid2mesh) if pats is not None else grid
        grid, mesh = self.encoder(grid, mesh, e_g2m, sg.grid2mesh_graph_src_indices, sg.grid2mesh_graph_dst_indices,
                                  grid_aug=grid_aug, agg_plan=plan("g2m", sg.grid2mesh_graph_dst_indices, n_mesh),
                                  edge_scatter_plans=[plan("g2m:src", sg.grid2mesh_graph_src_indices, grid_aug.shape[0]),
                                                      plan("g2m", sg.grid2mesh_graph_dst_indices, n_mesh), None] if use_plans else None)
        n_mesh_aug = n_mesh + (pats.mesh.num_halo_vertices if pats is not None else 0)
        mesh, _ = self.processor(mesh, e_mm, sg.mesh_graph_src_indices, sg.mesh_graph_dst_indices,
                                 exchange=(lambda n: halo(n, pats.mesh)) if pats is not None else None,
                                 agg_plan=plan("mesh", sg.mesh_graph_src_indices, n_mesh),
                                 edge_scatter_plans=[plan("mesh", sg.mesh_graph_src_indices, n_mesh),
                                                     plan("mesh:dst", sg.mesh_graph_dst_indices, n_mesh_aug), None] if use_plans else None)
        mesh_aug = halo(mesh, pats.mesh2grid) if pats is not None else mesh
        x = self.decoder(e_m2g, grid, mesh, sg.mesh2grid_graph_src_indices, sg.mesh2grid_graph_dst_indices,
                         mesh_aug=mesh_aug, agg_plan=plan("m2g", sg.mesh2grid_graph_dst_indices, n_grid),
                         edge_scatter_plans=[plan("m2g:src", sg.mesh2grid_graph_src_indices, mesh_aug.shape[0]),
                                             plan("m2g", sg.mesh2grid_graph_dst_indices, n_grid), None] if use_plans else None)
        out = self.final_prediction(x)
        return x_in + out[:, : x_in.shape[1]].to(dt)
