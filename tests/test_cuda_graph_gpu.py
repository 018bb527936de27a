"""Capturable Adam and whole-step CUDA graph replay vs the eager step (1 GPU)."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(V=4000, deg=8, F=24, H=64, C=7, seed=0):
    from dgraph_b200 import Communicator
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern
    from tests._graphs import random_graph

    dev = torch.device("cuda")
    Communicator._is_initialized = False
    comm = Communicator.init_process_group("single")
    e = random_graph(V, deg, seed=seed, symmetric=True).to(dev)
    pat = build_communication_pattern(e, torch.zeros(V, dtype=torch.long, device=dev), 0, 1)
    pat.kernel_plan(dev)
    torch.manual_seed(seed)
    model = CommAwareGCN(F, H, C, HaloExchange(comm), comm, compute_dtype=torch.bfloat16).to(dev)
    x = torch.randn(V, F, device=dev).to(torch.bfloat16)
    y = torch.randint(0, C, (V,), device=dev)
    idx = torch.arange(0, V, 3, device=dev)
    return comm, model, pat, x, y, idx


def test_capturable_adam_matches_regular():
    from dgraph_b200.optim import FusedAdam

    torch.manual_seed(1)
    ws = [torch.randn(37, 5, device="cuda"), torch.randn(129, device="cuda")]
    pa = [torch.nn.Parameter(w.clone()) for w in ws]
    pb = [torch.nn.Parameter(w.clone()) for w in ws]
    oa = FusedAdam(pa, lr=1e-2, weight_decay=1e-3)
    ob = FusedAdam(pb, lr=1e-2, weight_decay=1e-3, capturable=True)
    for t in range(5):
        if t == 3:
            oa.param_groups[0]["lr"] = ob.param_groups[0]["lr"] = 5e-3
        for o, ps in ((oa, pa), (ob, pb)):
            o.zero_grad()
            sum(((p * (i + 1 + t)) ** 2).sum() for i, p in enumerate(ps)).backward()
            o.step()
    for a, b in zip(pa, pb):
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-7)
    assert int(ob.state_dict()["state"][0]["step"]) == 5


def test_graphed_train_step_matches_eager():
    from dgraph_b200.ops import _native as N
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.utils.cuda_graph import GraphedTrainStep

    comm, model_e, pat, x, y, idx = _setup()
    model_g = copy.deepcopy(model_e)
    crit = torch.nn.CrossEntropyLoss()

    def fl(model):
        return lambda xx, yy: crit(model(xx, pat).index_select(0, idx).float(), yy.index_select(0, idx))

    oe = FusedAdam(model_e.parameters(), lr=1e-2)
    og = FusedAdam(model_g.parameters(), lr=1e-2, capturable=True)
    warm = 2
    step = GraphedTrainStep(fl(model_g), og, (x, y), warmup=warm)
    assert step.launches_per_replay > 10
    eager = []
    for _ in range(warm + 4):  # the capture pass records but does not execute, so it is not a step
        oe.zero_grad()
        loss = fl(model_e)(x, y)
        loss.backward()
        oe.step()
        eager.append(float(loss))
    l0 = N.launches()
    graphed = [float(step(x, y)) for _ in range(4)]
    assert N.launches() - l0 == 4 * step.launches_per_replay
    # same data, same init: bf16 kernels are deterministic except for atomics in the loss gather backward
    for a, b in zip(eager[warm:warm + 4], graphed):
        assert abs(a - b) <= 2e-2 * max(1.0, abs(a)), (eager, graphed)
    assert graphed[-1] < graphed[0]
    x2 = torch.randn_like(x.float()).to(torch.bfloat16)
    l_new = float(step(x2, y))  # new inputs flow through the static buffers
    assert abs(l_new - graphed[-1]) > 1e-4
    comm.destroy()
