"""Cross-check of the plan builders against the reference's own pure-PyTorch helper functions, used as an oracle when the
reference tree is present (``/root/reference`` or the installed copy under ``baseline/_ref``).  The functions are loaded by file
path so they do not collide with the ``DGraph`` compatibility package of this repository.  Only order-independent facts are
compared (the reference numbers halo vertices by global id, this framework in receive order: SURVEY Q8)."""
import importlib.util
import os

import pytest
import torch

from dgraph_b200.parallel import build_communication_pattern
from tests._graphs import random_graph

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANDIDATES = ["/root/reference/DGraph/distributed/commInfo.py",
              os.path.join(ROOT, "baseline", "_ref", "DGraph", "distributed", "commInfo.py")]


def _ref():
    for p in CANDIDATES:
        if os.path.exists(p):
            spec = importlib.util.spec_from_file_location("_ref_commInfo", p)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
    pytest.skip("reference tree not available")


@pytest.mark.parametrize("W,seed", [(2, 0), (3, 1), (4, 2)])
def test_plan_pieces_agree_with_reference_helpers(W, seed):
    ref = _ref()
    V = 16 * W
    e = random_graph(V, 5, seed=seed, symmetric=True)
    part = (torch.arange(V) * W // V).clamp(max=W - 1)  # rank-contiguous, as the reference requires
    for r in range(W):
        pat = build_communication_pattern(e, part, r, W)
        local = ref.compute_local_vertices(part, r)
        assert torch.equal(pat.local_vertices_global, local)
        halo = ref.compute_halo_vertices(e, part, r)
        assert torch.equal(torch.sort(pat.halo_vertices_global).values, halo)  # same set, different numbering
        # local edges, mapped back to global ids: same multiset as the edges whose first endpoint is owned by r
        aug = torch.cat([pat.local_vertices_global, pat.halo_vertices_global])
        mine = e[part[e[:, 0]] == r]
        got = torch.stack([pat.local_vertices_global[pat.local_edge_list[:, 0]], aug[pat.local_edge_list[:, 1]]], 1)
        key = lambda t: torch.sort(t[:, 0] * V + t[:, 1]).values  # noqa: E731
        assert torch.equal(key(got), key(mine))
        ref_le = ref.compute_local_edge_list(e, part, local, halo, r)
        ref_aug = torch.cat([local, halo])
        ref_got = torch.stack([local[ref_le[:, 0]], ref_aug[ref_le[:, 1]]], 1)
        assert torch.equal(key(ref_got), key(got))
        # boundary vertices: per destination rank the same set of global ids is sent
        s_idx, s_off = ref.compute_boundary_vertices(e, part, local, r, W)
        for q in range(W):
            a = set(local[s_idx[s_off[q]:s_off[q + 1]]].tolist())
            b = set(pat.local_vertices_global[pat.send_local_idx[pat.send_offset[q]:pat.send_offset[q + 1]]].tolist())
            assert a == b, (r, q)
        assert torch.equal(s_off.cpu().long(), pat.send_offset.cpu().long())


def test_reference_gcn_model_runs_unmodified_on_this_framework():
    """Drop-in check: the reference's own ``experiments/OGB/GCN.py`` (loaded by path, unmodified) imports ``DGraph`` -- here the
    compatibility package of this repository -- and, with the same weights, produces the same logits and gradients as
    ``dgraph_b200.models.gcn.CommAwareGCN`` on a 3-rank emulated world and on a single process."""
    path = "/root/reference/experiments/OGB/GCN.py"
    if not os.path.exists(path):
        path = os.path.join(ROOT, "baseline", "_ref", "experiments", "OGB", "GCN.py")
    if not os.path.exists(path):
        pytest.skip("reference tree not available")
    spec = importlib.util.spec_from_file_location("_ref_GCN", path)
    ref_gcn = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_gcn)

    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.parallel import HaloExchange
    from dgraph_b200.parallel.emulator import run_world

    V, F, H, C, W = 48, 6, 16, 5, 3
    e = random_graph(V, 5, seed=4, symmetric=True)
    xg = torch.randn(V, F, dtype=torch.float64, generator=torch.Generator().manual_seed(0))
    torch.manual_seed(1)
    proto = ref_gcn.CommAwareGCN(F, H, C, None, None).double()
    state = {k: v.clone() for k, v in proto.state_dict().items()}

    def job(comm):
        r, w = comm.get_rank(), comm.get_world_size()
        pat = build_communication_pattern(e, (torch.arange(V) * w // V).clamp(max=w - 1), r, w)
        x = xg[pat.local_vertices_global]
        outs = []
        for cls in (ref_gcn.CommAwareGCN, CommAwareGCN):
            m = cls(F, H, C, HaloExchange(comm), comm).double()
            m.load_state_dict(state)  # same parameter names in both implementations
            y = m(x, pat)
            (y ** 2).sum().backward()
            g = torch.cat([p.grad.reshape(-1) for p in m.parameters()])
            comm.all_reduce(g)
            outs.append((y.detach(), g))
        return pat.local_vertices_global, outs

    single = run_world(1, job)[0]
    for rows, ((y_ref, g_ref), (y_ours, g_ours)) in run_world(W, job):
        assert torch.allclose(y_ref, y_ours, atol=1e-9) and torch.allclose(g_ref, g_ours, atol=1e-8)
        assert torch.allclose(y_ours, single[1][1][0][rows], atol=1e-9)
        assert torch.allclose(g_ours, single[1][1][1], atol=1e-8)


def test_reference_gat_layer_runs_unmodified_on_this_framework():
    """The reference's ``CommAwareGAT`` (``experiments/OGB-LSC/RGAT.py``, loaded by path) drives ``comm.gather / scatter`` with
    ``comm_plan=`` through this repository's ``DGraph`` package on a 3-rank emulated world; with shared weights it must agree
    with ``dgraph_b200.models.rgat.CommAwareGAT`` (forward and parameter gradients)."""
    import sys

    d = "/root/reference/experiments/OGB-LSC"
    if not os.path.exists(os.path.join(d, "RGAT.py")):
        pytest.skip("reference tree not available")
    saved = list(sys.path)
    mods = set(sys.modules)
    try:
        sys.path.insert(0, d)  # its sibling modules (distributed_layers, CacheGenerator)
        spec = importlib.util.spec_from_file_location("_ref_RGAT", os.path.join(d, "RGAT.py"))
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
    finally:
        sys.path[:] = saved
        for m in set(sys.modules) - mods:
            if m in ("distributed_layers", "CacheGenerator", "config"):
                sys.modules.pop(m, None)

    from dgraph_b200.models.rgat import CommAwareGAT
    from dgraph_b200.parallel.comm_plan import COO_to_NCCLEdgeConditionedCommPlan
    from dgraph_b200.parallel.emulator import run_world

    W, V, F, H = 3, 36, 5, 8
    e = random_graph(V, 4, seed=7, symmetric=False)
    off = torch.tensor([V * r // W for r in range(W + 1)])
    xg = torch.randn(V, F, dtype=torch.float64, generator=torch.Generator().manual_seed(0))
    torch.manual_seed(3)
    proto = ref.CommAwareGAT(F, H, None, residual=True).double()
    state = {k: v.clone() for k, v in proto.state_dict().items()}

    def job(comm):
        r = comm.get_rank()
        src, dst = e[:, 0], e[:, 1]
        mine = torch.nonzero((src >= off[r]) & (src < off[r + 1]), as_tuple=True)[0]
        # edges live with their source; messages flow source -> destination (the layer's "dest" plan indexes dst)
        plan = COO_to_NCCLEdgeConditionedCommPlan(r, W, dst, src, mine, off, None, exchange=comm.engine.exchange_ids)
        x = xg[off[r]:off[r + 1]].unsqueeze(0)
        res = []
        for cls in (ref.CommAwareGAT, CommAwareGAT):
            m = cls(F, H, comm, residual=True).double()
            m.load_state_dict(state)
            y = m(x, comm_plan=plan)
            (y ** 2).sum().backward()
            g = torch.cat([p.grad.reshape(-1) for p in m.parameters()])
            comm.all_reduce(g)
            res.append((y.detach(), g))
        return res

    for (y_ref, g_ref), (y_ours, g_ours) in run_world(W, job):
        assert y_ref.shape == y_ours.shape
        assert torch.allclose(y_ref, y_ours, atol=1e-9) and torch.allclose(g_ref, g_ours, atol=1e-8)


def test_reference_single_process_test_files_pass_against_this_package(tmp_path):
    """The reference's own ``tests/test_comm_info.py`` (single-process cases) and ``tests/test_distributed_graph.py`` are run,
    unmodified, with this repository first on ``PYTHONPATH`` -- so ``import DGraph`` resolves to the compatibility package.  The
    eleven cases behind the ``dist_setup`` fixture need ``torchrun`` + NCCL on two GPUs and are deselected."""
    import subprocess
    import sys

    tdir = "/root/reference/tests"
    if not os.path.exists(os.path.join(tdir, "test_comm_info.py")):
        pytest.skip("reference tree not available")
    needs_nccl = ["test_compute_comm_map_correct_values", "test_compute_comm_map_row_matches_local_send_counts",
                  "test_compute_comm_map_diagonal_zero", "test_build_communication_pattern", "test_communication_pattern_invariants",
                  "test_communication_pattern_send_recv_symmetry", "test_hetero_halo_and_boundary_distributed"]
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(tdir, "test_comm_info.py"),
                        os.path.join(tdir, "test_distributed_graph.py"), "-q", "-p", "no:cacheprovider", "--rootdir", str(tmp_path),
                        "-k", " and ".join(f"not {n}" for n in needs_nccl)],
                       capture_output=True, text=True, cwd=str(tmp_path), env=env, timeout=600)
    tail = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-500:]
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
    assert "passed" in tail and "failed" not in tail and "error" not in tail, tail
    assert int(tail.split(" passed")[0].split()[-1]) >= 39, tail


@pytest.mark.timeout(300)
def test_reference_mpi_backend_tests_pass_against_this_package(tmp_path):
    """The reference's ``tests/test_mpi_backend.py`` (written for ``mpirun -n 2``: init, one-sided gather, one-sided scatter through
    ``Communicator.init_process_group("mpi")``), unmodified, as two CPU ranks started by ``tools/run_ranks.py``.  On a CPU box
    ``"mpi"`` is the shared-memory window engine (``parallel/shm_engine.py``), so the reference's cases exercise the same
    fence / get / accumulate epochs its mpi4py engine would."""
    import subprocess
    import sys

    from dgraph_b200.parallel import shm_engine

    f = "/root/reference/tests/test_mpi_backend.py"
    if not os.path.exists(f):
        pytest.skip("reference tree not available")
    if not shm_engine.available():
        pytest.skip("host library cannot be built (no C++ compiler)")
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_USE_AGENT_STORE"):
        env.pop(k, None)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "run_ranks.py"), "-n", "2", "--timeout", "240", "--", sys.executable,
                        "-m", "pytest", f, "-q", "-p", "no:cacheprovider", "--rootdir", str(tmp_path)],
                       capture_output=True, text=True, cwd=str(tmp_path), env=env, timeout=280)
    assert r.returncode == 0 and "3 passed" in r.stdout, r.stdout[-3000:] + r.stderr[-1000:]


@pytest.mark.gpu
def test_reference_local_kernel_tests_pass_against_this_package_gpu(tmp_path):
    """The reference's ``tests/test_local_kernels.py`` (its 1-GPU test of the ``torch_local`` extension), unmodified, with
    ``DGraph.torch_local`` resolving to this repository's launchers (``csrc/rank_local.cu``)."""
    import subprocess
    import sys

    # the GPU boxes have no /root/reference: fall back to the unmodified copy that tools/install_reference.sh keeps next to the
    # reference install (baseline/_ref travels to the box)
    cands = ["/root/reference/tests/test_local_kernels.py", os.path.join(ROOT, "baseline", "_ref", "tests", "test_local_kernels.py")]
    f = next((c for c in cands if os.path.exists(c)), None)
    if f is None:
        pytest.skip("reference tree not available")
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, "-m", "pytest", f, "-q", "-p", "no:cacheprovider", "--rootdir", str(tmp_path)],
                       capture_output=True, text=True, cwd=str(tmp_path), env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]


def test_every_public_name_of_the_reference_library_resolves():
    """Every top-level function / class of every module under the reference's ``DGraph/`` package exists under the same module
    path here (private MPI window helpers excepted: there is no MPI in this framework)."""
    import ast
    import importlib

    ref_root = "/root/reference/DGraph"
    if not os.path.isdir(ref_root):
        pytest.skip("reference tree not available")
    allowed_missing = {"_mpi_vector_get", "_mpi_vector_accumulate", "_mpi_gather_impl", "_mpi_scatter_add_impl"}
    problems = {}
    for d, _, files in os.walk(ref_root):
        for f in files:
            if not f.endswith(".py"):
                continue
            path = os.path.join(d, f)
            rel = os.path.relpath(path, "/root/reference")[:-3].replace(os.sep, ".")
            rel = rel[:-9] if rel.endswith(".__init__") else rel
            names = [n.name for n in ast.parse(open(path).read()).body
                     if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and not n.name.startswith("__")]
            try:
                mod = importlib.import_module(rel)
            except Exception as e:  # noqa: BLE001
                problems[rel] = repr(e)
                continue
            miss = [n for n in names if not hasattr(mod, n) and n not in allowed_missing]
            if miss:
                problems[rel] = miss
    assert not problems, problems


def test_every_method_of_the_reference_library_classes_exists():
    """Method-level version of the name check: every method of every class of the reference's ``DGraph/`` package exists on the
    class of the same name here (``forward`` / ``backward`` of the two legacy autograd functions excepted: they are plain
    callables with ``apply`` here)."""
    import ast
    import importlib

    ref_root = "/root/reference/DGraph"
    if not os.path.isdir(ref_root):
        pytest.skip("reference tree not available")
    problems = {}
    for d, _, files in os.walk(ref_root):
        for f in files:
            if not f.endswith(".py"):
                continue
            path = os.path.join(d, f)
            rel = os.path.relpath(path, "/root/reference")[:-3].replace(os.sep, ".")
            rel = rel[:-9] if rel.endswith(".__init__") else rel
            try:
                mod = importlib.import_module(rel)
            except Exception:  # noqa: BLE001 - module-level resolution is the previous test's job
                continue
            for node in ast.parse(open(path).read()).body:
                if not isinstance(node, ast.ClassDef) or not hasattr(mod, node.name):
                    continue
                cls = getattr(mod, node.name)
                meths = [m.name for m in node.body if isinstance(m, ast.FunctionDef) and not m.name.startswith("__")]
                miss = [m for m in meths if not hasattr(cls, m)]
                if node.name in ("GatherFunction", "ScatterFunction"):
                    miss = [m for m in miss if m not in ("forward", "backward")]
                if miss:
                    problems[f"{rel}.{node.name}"] = miss
    assert not problems, problems


def _load_by_path(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_graphcast_mesh_and_spatial_helpers_agree_with_reference():
    """``experiments/GraphCast/data_utils/{icosahedral_mesh,spatial_utils}.py`` of this repository against the reference's files
    of the same name: same vertex sets per refinement level (vertex numbering differs, so sets are compared), same multimesh
    edge set, same spherical-geometry functions."""
    import numpy as np
    base = "/root/reference/experiments/GraphCast/data_utils/"
    if not os.path.exists(base):
        base = os.path.join(ROOT, "baseline", "_ref", "experiments", "GraphCast", "data_utils") + "/"
    if not os.path.exists(base + "icosahedral_mesh.py"):
        pytest.skip("reference tree not available")
    pytest.importorskip("scipy")
    ri, rs = _load_by_path(base + "icosahedral_mesh.py", "_ref_ico"), _load_by_path(base + "spatial_utils.py", "_ref_sp")
    mine = os.path.join(ROOT, "experiments", "GraphCast", "data_utils") + "/"
    oi, osp = _load_by_path(mine + "icosahedral_mesh.py", "_our_ico"), _load_by_path(mine + "spatial_utils.py", "_our_sp")

    def vkey(v):  # order-free fingerprint of a vertex set
        q = np.round(np.asarray(v, dtype=np.float64) * 1e4).astype(np.int64)
        return sorted(map(tuple, q))

    a, b = ri.get_hierarchy_of_triangular_meshes_for_sphere(3), oi.get_hierarchy_of_triangular_meshes_for_sphere(3)
    assert len(a) == len(b) == 4
    for x, y in zip(a, b):
        assert x.vertices.shape == y.vertices.shape and x.faces.shape == y.faces.shape
        assert vkey(x.vertices) == vkey(y.vertices)
        # faces: same set of corner-coordinate triples, all outward-oriented
        cx, cy = rs.get_face_centroids(x.vertices, x.faces), osp.get_face_centroids(y.vertices, y.faces)
        assert vkey(cx) == vkey(cy)
        v, f = np.asarray(y.vertices, dtype=np.float64), np.asarray(y.faces)
        n = np.cross(v[f[:, 1]] - v[f[:, 0]], v[f[:, 2]] - v[f[:, 0]])
        assert ((n * v[f].mean(1)).sum(1) > 0).all()
    ma, mb = ri.merge_meshes(a), oi.merge_meshes(b)
    (sa, da), (sb, db) = ri.faces_to_edges(ma.faces), oi.faces_to_edges(mb.faces)

    def ekey(m, s, d):
        q = np.round(np.asarray(m.vertices, dtype=np.float64) * 1e4).astype(np.int64)
        return sorted(map(tuple, np.concatenate([q[s], q[d]], axis=1)))

    assert ekey(ma, sa, da) == ekey(mb, sb, db)
    assert abs(rs.max_edge_length(ma.vertices, sa, da) - osp.max_edge_length(mb.vertices, sb, db)) < 1e-6

    g = torch.Generator().manual_seed(0)
    ll = torch.rand(64, 2, generator=g) * torch.tensor([180.0, 360.0]) - torch.tensor([90.0, 180.0])
    th = torch.rand(64, generator=g) * 6 - 3
    for unit in ("deg", "rad"):
        q = ll if unit == "deg" else ll * (3.141592653589793 / 180)
        xa, xb = rs.latlon2xyz(q, unit=unit), osp.latlon2xyz(q, unit=unit)
        torch.testing.assert_close(xa, xb)
        torch.testing.assert_close(rs.xyz2latlon(xa, unit=unit), osp.xyz2latlon(xb, unit=unit))
    # rotations are compared in radians only: with unit="deg" the reference rescales the *points* instead of converting the
    # angle (spatial_utils.py:107-108); here degrees mean degrees
    xa = rs.latlon2xyz(ll)
    for ax in "xyz":
        torch.testing.assert_close(rs.geospatial_rotation(xa, th, ax), osp.geospatial_rotation(xa, th, ax))
        torch.testing.assert_close(osp.geospatial_rotation(xa, th * (180 / 3.141592653589793), ax, unit="deg"),
                                   osp.geospatial_rotation(xa, th, ax), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(rs.polar_angle(th), osp.polar_angle(th))
    torch.testing.assert_close(rs.azimuthal_angle(th), osp.azimuthal_angle(th))
    assert abs(rs.deg2rad(33.0) - osp.deg2rad(33.0)) < 1e-12 and abs(rs.rad2deg(0.7) - osp.rad2deg(0.7)) < 1e-12
    # the incremental child-vertex builder hands out one index per parent edge, whatever the order of the two endpoints
    cb = oi._ChildVerticesBuilder(b[0].vertices)
    i = cb.get_new_child_vertex_index((0, 1))
    assert cb.get_new_child_vertex_index((1, 0)) == i and cb.get_all_vertices().shape == (13, 3)


def test_every_name_of_the_reference_experiment_modules_resolves():
    """Same check as for the library, for ``experiments/``: every top-level function / class of every reference experiment module,
    and every method of those classes, exists under the same relative path and name here.  Exceptions are listed with reasons."""
    import ast
    import importlib
    import sys

    ref = "/root/reference/experiments"
    if not os.path.isdir(ref):
        ref = os.path.join(ROOT, "baseline", "_ref", "experiments")
    if not os.path.isdir(os.path.join(ref, "GraphCast")):
        pytest.skip("reference experiments tree not available")
    ours = os.path.join(ROOT, "experiments")
    skip_files = {
        # environment probes for NVSHMEM / MPI installs: there is no such dependency here (PeerMemory-Enabled-DGRAPH instead)
        "NVSHMEM-Enabled-DGRAPH/PreInstallCheck.py", "NVSHMEM-Enabled-DGRAPH/NVSHMEM_init.py",
    }
    skip_names = {
        ("GraphCast/dataset.py", "test_synthetic_weather_dataset"),  # an ad-hoc test function inside the module; see tests/
        ("OGB-LSC/lsc_datasets/distributed_graph_dataset.py", "DistributedHeteroGraphDataset.num_classes"),  # instance attribute here
    }

    def load(path):
        d = os.path.dirname(path)
        extra = [d, os.path.dirname(d)]
        for p in extra:
            sys.path.insert(0, p)
        try:
            if os.path.basename(d) == "lsc_datasets":
                return importlib.import_module("lsc_datasets." + os.path.basename(path)[:-3])
            return _load_by_path(path, "_expmod_" + os.path.relpath(path, ours).replace(os.sep, "_").replace("-", "_")[:-3])
        finally:
            for p in extra:
                sys.path.remove(p)

    missing = []
    for root, _, files in os.walk(ref):
        for f in sorted(files):
            rel = os.path.relpath(os.path.join(root, f), ref)
            if not f.endswith(".py") or f == "__init__.py" or "/tests/" in "/" + rel or rel in skip_files:
                continue
            tree = ast.parse(open(os.path.join(root, f)).read())
            mine = os.path.join(ours, rel)
            assert os.path.exists(mine), f"experiments/{rel} does not exist"
            mod = load(mine)
            for node in tree.body:
                if isinstance(node, (ast.FunctionDef, ast.ClassDef)):
                    if (rel, node.name) in skip_names:
                        continue
                    if not hasattr(mod, node.name):
                        missing.append(f"{rel}: {node.name}")
                        continue
                    if isinstance(node, ast.ClassDef):
                        cls = getattr(mod, node.name)
                        for m in (b for b in node.body if isinstance(b, ast.FunctionDef)):
                            key = (rel, f"{node.name}.{m.name}")
                            if key not in skip_names and not hasattr(cls, m.name):
                                missing.append(f"{rel}: {node.name}.{m.name}")
    assert not missing, "; ".join(missing)


def test_positional_signatures_of_the_reference_library_are_accepted():
    """Argument-level parity: for every function / method of the reference's ``DGraph`` package, a call that passes the
    reference's parameters positionally or by the reference's keyword names binds here too (same position, or the name is
    accepted).  Autograd ``Function.forward / backward`` internals and the one-sided engines' tensor-argument names are exempt
    (listed below with reasons)."""
    import ast
    import copy
    import importlib
    import inspect

    ref = "/root/reference/DGraph"
    if not os.path.isdir(ref):
        ref = os.path.join(ROOT, "baseline", "_ref", "DGraph")
    if not os.path.isdir(ref):
        pytest.skip("reference tree not available")
    exempt = (
        ".forward", ".backward",  # autograd Function internals: called by torch with positional tensors only
        # engines: the first tensor argument is positional in every caller; its keyword spellings (local_send_tensor / send_tensor /
        # input_tensor) and num_output_rows / rank_mapping are accepted through **kwargs (tests/test_api_surface.py)
        "NCCLBackendEngine.scatter", "NCCLBackendEngine.gather", "MPIBackendEngine.scatter", "MPIBackendEngine.gather",
        "NVSHMEMBackendEngine.scatter", "NVSHMEMBackendEngine.gather",
        "DGraph.utils.largest_split",  # the reference has two definitions (the shadowed DGraph/utils.py names them differently)
        "MPIBackendEngine.to_global_rank",  # needs the engine instance here: there is no class-level replica state
    )
    problems = []

    def check(name, node, obj):
        if name.endswith(exempt):
            return
        try:
            sig = inspect.signature(obj)
        except (TypeError, ValueError):
            return
        ps = list(sig.parameters.values())
        pos = [p.name for p in ps if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]
        names = {p.name for p in ps}
        var_pos = any(p.kind == p.VAR_POSITIONAL for p in ps)
        var_kw = any(p.kind == p.VAR_KEYWORD for p in ps)
        a = node.args
        for i, arg in enumerate(a.posonlyargs + a.args):
            if i < len(pos):
                if pos[i] != arg.arg:
                    problems.append(f"{name}: parameter {i} is '{pos[i]}' here, '{arg.arg}' in the reference")
            elif not var_pos and not (arg.arg in names or var_kw):
                problems.append(f"{name}: no place for parameter {i} '{arg.arg}'")
        for arg in a.kwonlyargs:
            if arg.arg not in names and not var_kw:
                problems.append(f"{name}: keyword '{arg.arg}' not accepted")

    for root, _, files in os.walk(ref):
        for f in sorted(files):
            if not f.endswith(".py"):
                continue
            rel = os.path.relpath(os.path.join(root, f), os.path.dirname(ref))
            modname = rel[:-3].replace(os.sep, ".")
            modname = modname[:-9] if modname.endswith(".__init__") else modname
            try:
                mod = importlib.import_module(modname)
            except ImportError:
                continue
            tree = ast.parse(open(os.path.join(root, f)).read())
            for node in tree.body:
                if isinstance(node, ast.FunctionDef) and hasattr(mod, node.name):
                    check(f"{modname}.{node.name}", node, getattr(mod, node.name))
                elif isinstance(node, ast.ClassDef) and hasattr(mod, node.name):
                    cls = getattr(mod, node.name)
                    for m in (b for b in node.body if isinstance(b, ast.FunctionDef)):
                        if not hasattr(cls, m.name):
                            continue
                        obj = getattr(cls, m.name)
                        m2 = copy.deepcopy(m)
                        static = isinstance(inspect.getattr_static(cls, m.name), staticmethod)
                        if not static and m2.args.args and m2.args.args[0].arg in ("self", "cls", "ctx"):
                            m2.args.args = m2.args.args[1:]
                            try:
                                sig = inspect.signature(obj)
                            except (TypeError, ValueError):
                                continue
                            ps = list(sig.parameters.values())
                            if ps and ps[0].name in ("self", "cls", "ctx"):
                                shim = lambda *a, **k: None  # noqa: E731
                                shim.__signature__ = sig.replace(parameters=ps[1:])
                                obj = shim
                        check(f"{modname}.{node.name}.{m.name}", m2, obj)
    assert not problems, "\n".join(problems)


def test_positional_signatures_of_the_reference_experiment_modules_are_accepted():
    """Argument-level parity for ``experiments/``: a positional call written against a reference function / method binds to the
    parameter of the same name here.  Exemptions (with reasons) are listed; constructors with two calling forms are checked by
    ``tests/test_experiment_helpers.py`` instead."""
    import ast
    import copy
    import importlib
    import inspect
    import sys

    ref = "/root/reference/experiments"
    if not os.path.isdir(ref):
        ref = os.path.join(ROOT, "baseline", "_ref", "experiments")
    if not os.path.isdir(os.path.join(ref, "GraphCast")):
        pytest.skip("reference experiments tree not available")
    ours = os.path.join(ROOT, "experiments")
    exempt = {
        # private timing helpers of one script, rewritten around this package's static graph object
        "GraphCast/microbenchmark_graphcast.py:load_data", "GraphCast/microbenchmark_graphcast.py:time_mesh_edge_block",
        "OGB/main.py:_run_experiment",
        # legacy cache plumbing (three cache objects + three file names upstream; one relation at a time here)
        "OGB-LSC/CacheGenerator.py:get_cache",
        "OGB-LSC/lsc_datasets/distributed_graph_dataset.py:DistributedHeteroGraphDataset._save_comm_plans",
        "OGB-LSC/lsc_datasets/distributed_graph_dataset.py:DistributedHeteroGraphDataset._load_comm_plans",
        "OGB-LSC/lsc_datasets/distributed_graph_dataset.py:DistributedHeteroGraphDataset._generate_comm_plans",
        # autograd internals / static helper / legacy COO path with this package's keyword-only arguments
        "OGB-LSC/RGAT.py:CommAwareGAT._calc_attention_messages", "OGB-LSC/RGAT.py:CommAwareGAT._forward_coo",
        "OGB-LSC/distributed_layers.py:DistributedBN_Impl.forward", "OGB-LSC/distributed_layers.py:DistributedBN_Impl.backward",
        "OGB-LSC/distributed_layers.py:DistributedBN_with_Recompute.forward",
        "OGB-LSC/distributed_layers.py:DistributedBN_with_Recompute.backward",
        # two calling forms, dispatched on the type of the first argument (tests/test_experiment_helpers.py)
        "OGB-LSC/lsc_datasets/MAG240M_dataset.py:DGraph_MAG240M_Dataset.__init__",
        "OGB-LSC/lsc_datasets/synthetic_dataset.py:SyntheticHeterogeneousDataset.__init__",
        "OGB-LSC/lsc_datasets/distributed_graph_dataset.py:DistributedHeteroGraphDataset.__init__",
    }

    def load(path):
        d = os.path.dirname(path)
        extra = [d, os.path.dirname(d)]
        for p in extra:
            sys.path.insert(0, p)
        try:
            if os.path.basename(d) == "lsc_datasets":
                return importlib.import_module("lsc_datasets." + os.path.basename(path)[:-3])
            return _load_by_path(path, "_expsig_" + os.path.relpath(path, ours).replace(os.sep, "_").replace("-", "_")[:-3])
        finally:
            for p in extra:
                sys.path.remove(p)

    problems = []

    def check(name, node, obj):
        if name in exempt:
            return
        try:
            sig = inspect.signature(obj)
        except (TypeError, ValueError):
            return
        ps = list(sig.parameters.values())
        pos = [p.name for p in ps if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]
        names = {p.name for p in ps}
        var_pos = any(p.kind == p.VAR_POSITIONAL for p in ps)
        var_kw = any(p.kind == p.VAR_KEYWORD for p in ps)
        for i, arg in enumerate(node.args.posonlyargs + node.args.args):
            if i < len(pos):
                if pos[i] != arg.arg:
                    problems.append(f"{name}: parameter {i} is '{pos[i]}' here, '{arg.arg}' in the reference")
            elif not var_pos and not (arg.arg in names or var_kw):
                problems.append(f"{name}: no place for parameter {i} '{arg.arg}'")

    for root, _, files in os.walk(ref):
        for f in sorted(files):
            rel = os.path.relpath(os.path.join(root, f), ref)
            if not f.endswith(".py") or f == "__init__.py" or "/tests/" in "/" + rel or rel.startswith("NVSHMEM-Enabled"):
                continue
            mine = os.path.join(ours, rel)
            if not os.path.exists(mine):
                continue
            tree = ast.parse(open(os.path.join(root, f)).read())
            mod = load(mine)
            for node in tree.body:
                if isinstance(node, ast.FunctionDef) and hasattr(mod, node.name):
                    check(f"{rel}:{node.name}", node, getattr(mod, node.name))
                elif isinstance(node, ast.ClassDef) and hasattr(mod, node.name):
                    cls = getattr(mod, node.name)
                    for m in (b for b in node.body if isinstance(b, ast.FunctionDef)):
                        if not hasattr(cls, m.name):
                            continue
                        obj, m2 = getattr(cls, m.name), copy.deepcopy(m)
                        static = isinstance(inspect.getattr_static(cls, m.name), staticmethod)
                        if not static and m2.args.args and m2.args.args[0].arg in ("self", "cls", "ctx"):
                            m2.args.args = m2.args.args[1:]
                            try:
                                sig = inspect.signature(obj)
                            except (TypeError, ValueError):
                                continue
                            ps = list(sig.parameters.values())
                            if ps and ps[0].name in ("self", "cls", "ctx"):
                                shim = lambda *a, **k: None  # noqa: E731
                                shim.__signature__ = sig.replace(parameters=ps[1:])
                                obj = shim
                        check(f"{rel}:{node.name}.{m.name}", m2, obj)
    assert not problems, "\n".join(problems)
