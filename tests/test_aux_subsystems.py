"""Tracing hooks, plan validators, watchdog / heartbeat / health check (all CPU)."""
import json
import time

import pytest
import torch

from dgraph_b200.ops import _native as N
from dgraph_b200.parallel import build_communication_pattern
from dgraph_b200.parallel.kernel_plan import GraphKernelPlan
from dgraph_b200.utils import debug, tracing, watchdog
from tests._graphs import random_graph, partition


def test_kernel_trace_records_launch_hooks(tmp_path):
    with tracing.KernelTrace() as tr:
        N.check(0, "dgb_fake_a")
        N.check(0, "dgb_fake_b")
        N.check(0, "dgb_fake_a")
    assert [r["name"] for r in tr.records] == ["dgb_fake_a", "dgb_fake_b", "dgb_fake_a"]
    assert tr.summary()["dgb_fake_a"]["launches"] == 2
    assert not N._post_launch_hooks
    p = tmp_path / "t.json"
    tr.to_chrome_trace(str(p))
    assert len(json.load(open(p))["traceEvents"]) == 3
    with tracing.nvtx_range("x"):
        pass
    assert "aten::add" in tracing.profile_step(lambda: torch.ones(4) + 1)


def test_plan_validators_accept_good_and_reject_corrupt():
    e = random_graph(30, 4, seed=3, symmetric=True)
    part = partition(30, 3, "random")
    for r in range(3):
        pat = build_communication_pattern(e, part, r, 3)
        debug.validate_pattern(pat)
        kp = pat.kernel_plan("cpu")
        debug.validate_kernel_plan(kp)
    bad = GraphKernelPlan.from_edges(torch.tensor([[0, 1], [1, 0], [1, 2]]), 2, 3, device="cpu")
    debug.validate_kernel_plan(bad)
    bad.col = bad.col.clone()
    bad.col[0] = 7
    with pytest.raises(AssertionError, match="csr.col"):
        debug.validate_kernel_plan(bad)
    pat.send_local_idx = pat.send_local_idx.clone()
    if pat.send_local_idx.numel():
        pat.send_local_idx[0] = pat.num_local_vertices + 5
        with pytest.raises(AssertionError):
            debug.validate_pattern(pat)
    with pytest.raises(FloatingPointError, match="w"):
        debug.check_finite([("w", torch.tensor([1.0, float("nan")])), ("v", torch.ones(2))])


def test_watchdog_fires_only_without_beats():
    fired = []
    with watchdog.StepWatchdog(0.3, on_timeout=fired.append, poll_s=0.05) as wd:
        for _ in range(8):
            time.sleep(0.05)
            wd.beat()
        assert not fired
        time.sleep(0.6)
    assert fired and wd.fired


def test_heartbeat_and_health_check(tmp_path, monkeypatch):
    hb0, hb1 = watchdog.Heartbeat(str(tmp_path), 0, 2), watchdog.Heartbeat(str(tmp_path), 1, 2)
    hb0.beat(5)
    assert hb0.stale(60.0) == [1]
    hb1.beat(5)
    assert hb0.stale(60.0) == [] and hb1.read(0)[0] == 5
    from dgraph_b200.utils.dist_utils import SingleProcessDummyCommunicator

    comm = SingleProcessDummyCommunicator()
    assert watchdog.health_check(comm, torch.tensor(1.0))
    assert not watchdog.health_check(comm, torch.tensor(float("inf")))
    monkeypatch.setenv("DGRAPH_B200_FAULT_INJECT", "nan:rank=0:step=2")
    assert torch.isfinite(watchdog.inject_fault(0, 1, torch.tensor(1.0)))
    assert torch.isnan(watchdog.inject_fault(0, 2, torch.tensor(1.0)))


def test_rowmid_and_restricted_csc_match_brute_force():
    torch.manual_seed(0)
    n_local, n_halo, E = 40, 25, 400
    e = torch.unique(torch.stack([torch.randint(0, n_local, (E,)), torch.randint(0, n_local + n_halo, (E,))], 1), dim=0)
    kp = GraphKernelPlan.from_edges(e, n_local, n_local + n_halo, device="cpu", num_nbr_local=n_local)
    mid = kp.rowmid()
    for r in range(n_local):
        cols = kp.col[kp.rowptr[r]:kp.rowptr[r + 1]]
        assert bool((cols[: mid[r] - kp.rowptr[r]] < n_local).all()) and bool((cols[mid[r] - kp.rowptr[r]:] >= n_local).all())
    rows = torch.arange(0, n_local, 3)
    colptr, row, smap = kp.restricted_csc(rows)
    assert kp.restricted_csc(rows)[0] is colptr  # cached per index tensor
    member = set(rows.tolist())
    for j in range(n_local + n_halo):
        full = [(int(kp.row[k]), int(kp.csc_to_csr[k])) for k in range(kp.colptr[j], kp.colptr[j + 1])]
        want = [(i, s) for i, s in full if i in member]
        got = [(int(row[k]), int(smap[k])) for k in range(colptr[j], colptr[j + 1])]
        assert got == want
    rows[0] = 1  # modified in place: the cached plan must not be served for the new contents
    colptr2, row2, _ = kp.restricted_csc(rows)
    assert colptr2 is not colptr and set(row2.tolist()) <= set(rows.tolist()) and 0 not in set(row2.tolist())
    inv = kp.csr_to_csc()
    assert torch.equal(inv[kp.csc_to_csr.long()].long(), torch.arange(kp.num_edges))


def test_sparsity_hint_is_one_shot_and_identity_checked():
    from dgraph_b200.ops import _sparsity
    from dgraph_b200.ops.loss import masked_cross_entropy, masked_cross_entropy_ref

    g = torch.zeros(6, 3)
    idx = torch.tensor([1, 4])
    _sparsity.put(g, idx, torch.ones(2, 3))
    assert _sparsity.take(torch.zeros(6, 3)) is None  # another tensor: no match, and the slot is consumed
    assert _sparsity.take(g) is None
    _sparsity.put(g, idx, torch.ones(2, 3))
    g.add_(1.0)  # modified after the note was left: stale
    assert _sparsity.take(g) is None
    _sparsity.put(g, idx, torch.ones(2, 3))
    got = _sparsity.take(g)
    assert got is not None and got[0] is idx
    logits = torch.randn(9, 4, dtype=torch.float64, requires_grad=True)
    y = torch.randint(0, 4, (9,))
    sel = torch.tensor([0, 3, 8])
    a = masked_cross_entropy(logits, y, sel)  # CPU: reference path
    assert torch.allclose(a, masked_cross_entropy_ref(logits, y, sel))
    assert torch.allclose(masked_cross_entropy(logits, y), torch.nn.functional.cross_entropy(logits, y))


def test_packed_counter_arithmetic_of_the_aggregation_kernel():
    """Host re-statement of the borrow-free packed counters in ``csrc/segment_ops.cu`` (``add_relu_pair``): per edge the kernel
    does ``word -= mask`` with 0xFFFF per active 16-bit half, repairs the high half once per row with ``word += word << 16``,
    and an accumulate pass resumes from stored counts with ``word - (word << 16)``."""
    import random

    M = 0xFFFFFFFF
    rnd = random.Random(0)
    for _ in range(200):
        n1, n2 = rnd.randrange(0, 300), rnd.randrange(0, 300)
        acts = [(rnd.random() < 0.6, rnd.random() < 0.3) for _ in range(n1 + n2)]

        def run(word, seq):
            for lo, hi in seq:
                mk = (0xFFFF if lo else 0) | (0xFFFF0000 if hi else 0)
                word = (word - mk) & M
            return word

        def repair(word):
            return (word + (word << 16)) & M

        one = repair(run(0, acts))
        lo_true, hi_true = sum(a for a, _ in acts), sum(b for _, b in acts)
        assert one == ((hi_true << 16) | lo_true)
        first = repair(run(0, acts[:n1]))  # pass 1 stores true counts ...
        resumed = (first - (first << 16)) & M  # ... pass 2 converts them back before continuing
        assert repair(run(resumed, acts[n1:])) == one


def test_symmetric_slot_rings_bookkeeping():
    """Host-side slot logic of the peer-memory engine with a fake arena (no GPU): rings grow on demand, a tracked entry stays
    reserved exactly as long as the tensor handed out for it is alive (whoever holds it, e.g. autograd's saved tensors), and re-using
    a slot that peers read in the current epoch inserts a barrier first."""
    from dgraph_b200.parallel.peer_engine import PeerMemoryEngine

    class FakeArena:
        def __init__(self):
            self.slots, self.top = {}, 0

        def slot(self, name, nbytes):
            if name not in self.slots:
                self.slots[name] = self.top
                self.top += (nbytes + 255) // 256 * 256
            return self.slots[name]

        def tensor(self, off, shape, dtype, guard=None):
            t = torch.empty(shape, dtype=dtype)
            t._guard = guard  # the real arena puts the guard into the storage's owner
            return t

    eng = object.__new__(PeerMemoryEngine)
    eng._world, eng._rank, eng._epoch, eng._rings, eng._by_off, eng.arena = 2, 0, 0, {}, {}, FakeArena()
    barriers = []

    def fake_barrier():
        eng._epoch += 1
        barriers.append(eng._epoch)

    eng.barrier_on_stream = fake_barrier
    t1, o1 = eng.slot_tensor("halo", (4, 8), torch.float32, ring=True, track=True)
    t2, o2 = eng.slot_tensor("halo", (4, 8), torch.float32, ring=True, track=True)
    assert o1 != o2 and t1.shape == (4, 8)  # first buffer is still alive (saved for backward) -> a second entry is added
    saved = [t1]  # any holder counts, not only the producer's autograd node
    del t1
    _, o2b = eng.slot_tensor("halo", (4, 8), torch.float32, ring=True, track=True)
    assert o2b not in (o1, o2)
    saved.clear()  # last reference gone -> entry free again
    t3, o3 = eng.slot_tensor("halo", (4, 8), torch.float32, ring=True, track=True)
    assert o3 == o1  # released entry is re-used, the ring does not grow
    del t3, t2
    eng.new_step()
    _, o4 = eng.slot_tensor("halo", (4, 8), torch.float32, ring=True, track=False)
    assert o4 == o1 and not barriers
    eng.mark_read(o4)  # peers read it after the latest barrier ...
    _, o5 = eng.slot_tensor("halo", (4, 8), torch.float32, ring=True, track=False)
    assert o5 == o1 and barriers == [1]  # ... so overwriting it needs one more barrier, exactly once
    _, o6 = eng.slot_tensor("halo", (4, 8), torch.float32, ring=True, track=False)
    assert barriers == [1]
    _, p1 = eng.slot_tensor("plain", (16,), torch.bfloat16)
    _, p2 = eng.slot_tensor("plain", (16,), torch.bfloat16)
    assert p1 == p2 and p1 not in (o1, o2)  # non-ring keys are one persistent buffer
    _, q = eng.slot_tensor("halo", (8, 8), torch.float32, ring=True)
    assert q not in (o1, o2)  # another shape under the same key gets its own ring
