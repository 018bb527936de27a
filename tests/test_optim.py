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


def test_checkpoint_save_and_resume_reproduces_training(tmp_path):
    """Model file = plain state_dict (reference format, also loadable with a DDP ``module.`` prefix); the side file restores the
    fused optimizer's moments / step and the scheduler, so a resumed run continues bit-for-bit."""
    import torch
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.utils.checkpoint import load_checkpoint, save_checkpoint

    def make():
        torch.manual_seed(0)
        m = torch.nn.Sequential(torch.nn.Linear(6, 8), torch.nn.Tanh(), torch.nn.Linear(8, 3))
        o = FusedAdam(m.parameters(), lr=1e-2, weight_decay=1e-3)
        s = torch.optim.lr_scheduler.StepLR(o, step_size=2, gamma=0.5)
        return m, o, s

    x = torch.randn(32, 6, generator=torch.Generator().manual_seed(1))
    y = torch.randn(32, 3, generator=torch.Generator().manual_seed(2))

    def steps(m, o, s, n):
        out = []
        for _ in range(n):
            o.zero_grad()
            loss = ((m(x) - y) ** 2).mean()
            loss.backward()
            o.step()
            s.step()
            out.append(float(loss))
        return out

    m, o, s = make()
    steps(m, o, s, 3)
    path = str(tmp_path / "model_3.pth")
    save_checkpoint(path, m, o, s, step=3)
    want = steps(m, o, s, 3)
    plain = torch.load(path, weights_only=True)
    assert set(plain) == {"0.weight", "0.bias", "2.weight", "2.bias"}  # nothing but the state_dict in the model file
    m2, o2, s2 = make()
    assert load_checkpoint(path, m2, o2, s2) == 3
    assert steps(m2, o2, s2, 3) == want
    # a file written from a DDP-wrapped reference model carries "module." prefixes
    torch.save({"module." + k: v for k, v in plain.items()}, str(tmp_path / "ddp.pth"))
    m3, _, _ = make()
    load_checkpoint(str(tmp_path / "ddp.pth"), m3)
    assert all(torch.equal(a, b) for a, b in zip(m3.state_dict().values(), plain.values()))
