import pytest
import torch

from dgraph_b200.optim import FusedAdam


def _run(device):
    torch.manual_seed(0)
    a = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3)).to(device)
    b = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3)).to(device)
    b.load_state_dict(a.state_dict())
    oa = FusedAdam(a.parameters(), lr=1e-2, weight_decay=1e-3)
    ob = torch.optim.Adam(b.parameters(), lr=1e-2, weight_decay=1e-3)
    x = torch.randn(11, 7, device=device)
    for _ in range(5):
        for m, o in ((a, oa), (b, ob)):
            o.zero_grad()
            m(x).square().mean().backward()
            o.step()
    for pa, pb in zip(a.parameters(), b.parameters()):
        assert torch.allclose(pa, pb, atol=1e-6, rtol=1e-5)
    sd = oa.state_dict()
    ref = ob.state_dict()
    assert torch.allclose(sd["state"][0]["exp_avg"], ref["state"][0]["exp_avg"], atol=1e-6)
    oc = FusedAdam(b.parameters(), lr=1e-2)
    oc.load_state_dict(sd)
    assert oc._step == 5


def test_fused_adam_cpu_matches_torch():
    _run("cpu")


@pytest.mark.gpu
def test_fused_adam_gpu_matches_torch():
    _run("cuda")
