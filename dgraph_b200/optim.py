"""Fused flat-arena Adam (+ data-parallel gradient all-reduce).

Replaces the reference's ``torch.optim.Adam`` + DDP pair (``experiments/OGB/main.py:112-114``): parameters and
gradients are re-homed into two contiguous fp32 buffers, so

* the gradient all-reduce of data-parallel training is ONE collective over the flat gradient buffer (the role of
  DDP's bucketed all-reduce), and
* the optimizer step is ONE sm_100a kernel (``csrc/adam.cu``) that also folds in the 1/world gradient scale.

State-dict layout follows ``torch.optim.Adam`` (``state[p] = {step, exp_avg, exp_avg_sq}``) so checkpoints
round-trip with stock PyTorch.
"""
from __future__ import annotations

import math
from typing import Iterable, Optional

import torch

from dgraph_b200.ops import _native as N
from dgraph_b200.ops import _wcache


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params: Iterable, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, comm=None, grad_average: bool = True, capturable: bool = False):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        assert len(self.param_groups) == 1, "FusedAdam keeps one flat arena (single param group)"
        self.comm = comm
        self.grad_average = grad_average
        self._step = 0
        self.capturable = capturable  # step counter / lr in device memory: the step can live inside a CUDA graph
        ps = [p for p in self.param_groups[0]["params"] if p.requires_grad]
        assert all(p.dtype == torch.float32 for p in ps), "master parameters are fp32"
        dev = ps[0].device
        sizes = [(p.numel() + 3) // 4 * 4 for p in ps]  # keep every view 16-byte aligned
        total = sum(sizes)
        self.flat_p = torch.zeros(total, dtype=torch.float32, device=dev)
        # peer-memory engine: gradients live in the symmetric arena so the Adam kernel can sum all ranks' buffers itself
        self._peer = None
        eng = getattr(comm, "engine", None)
        if eng is not None and hasattr(eng, "adam_allreduce") and self._global_world() > 1 and dev.type == "cuda":
            self._peer = eng
            self.flat_g, self._g_off = eng.symmetric_flat("adam_grad", total)
            self.flat_g = self.flat_g[:total]
            self.flat_g.zero_()
        else:
            self.flat_g = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_m = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_v = torch.zeros(total, dtype=torch.float32, device=dev)
        if capturable:
            assert dev.type == "cuda", "capturable FusedAdam is a CUDA-graph feature"
            self._dev_step = torch.zeros(1, dtype=torch.int32, device=dev)
            self._dev_state = torch.zeros(4, dtype=torch.float32, device=dev)
            self._dev_lr = None
            self.sync_hyper()
        self._views = []
        o = 0
        for p, s in zip(ps, sizes):
            n = p.numel()
            self.flat_p[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.flat_p[o:o + n].view(p.shape)
            p.grad = self.flat_g[o:o + n].view(p.shape)
            self._views.append((p, o, n))
            o += s
        self._params = ps

    def sync_hyper(self):
        """Capturable mode: push the (possibly scheduler-updated) learning rate to device memory.  Call it outside the
        captured region, before each replay; a no-op when the value did not change."""
        lr = float(self.param_groups[0]["lr"])
        if self._dev_lr != lr:
            self._dev_state[0:1].fill_(lr)
            self._dev_lr = lr

    def _global_world(self) -> int:
        c = self.comm
        return c.get_global_world_size() if hasattr(c, "get_global_world_size") else c.get_world_size()

    # gradients live in the flat buffer: zeroing is one memset and the views stay attached
    def zero_grad(self, set_to_none: bool = False):
        self.flat_g.zero_()
        for p, o, n in self._views:
            if p.grad is None or p.grad.data_ptr() != self.flat_g.data_ptr() + 4 * o:
                p.grad = self.flat_g[o:o + n].view(p.shape)

    def sync_grads(self):
        """Data-parallel all-reduce of the flat gradient (SUM); averaging is folded into the Adam kernel."""
        if self.comm is not None and self._global_world() > 1:
            self.comm.all_reduce(self.flat_g, scope="world") if hasattr(self.comm, "get_num_replicas") else self.comm.all_reduce(self.flat_g)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        _wcache.bump()  # the kernels below rewrite the parameters behind PyTorch's back: packed bf16 copies are stale
        for p, o, n in self._views:  # a grad that autograd re-allocated (first backward after set_to_none) is folded in
            if p.grad is not None and p.grad.data_ptr() != self.flat_g.data_ptr() + 4 * o:
                self.flat_g[o:o + n].add_(p.grad.reshape(-1))
                p.grad = self.flat_g[o:o + n].view(p.shape)
        self._step += 1
        g = self.param_groups[0]
        W = self._global_world() if (self.comm is not None and self.grad_average) else 1
        lr, (b1, b2), eps, wd = g["lr"], g["betas"], g["eps"], g["weight_decay"]
        if self.capturable:
            if not torch.cuda.is_current_stream_capturing():
                self.sync_hyper()
            if self._peer is not None:  # gradient sum over NVLink + Adam, step / lr read from device memory
                self._peer.adam_allreduce_capturable(self.flat_p, self._g_off, self.flat_m, self.flat_v, self.flat_p.numel(),
                                                     b1, b2, eps, wd, 1.0 / W, self._dev_step, self._dev_state)
                return loss
            self.sync_grads()
            N.check(N.lib().dgb_adam_flat_capturable(N.ptr(self.flat_p), N.ptr(self.flat_g), N.ptr(self.flat_m),
                                                     N.ptr(self.flat_v), N.ptr(None), N.i64(self.flat_p.numel()),
                                                     N.f32(b1), N.f32(b2), N.f32(eps), N.f32(wd), N.f32(1.0 / W),
                                                     N.ptr(self._dev_step), N.ptr(self._dev_state), N.stream()),
                    "dgb_adam_flat_capturable")
            N.count_launch(2)
            return loss
        if self._peer is not None:  # ONE kernel: sum of all ranks' gradients over NVLink + Adam
            self._peer.adam_allreduce(self.flat_p, self._g_off, self.flat_m, self.flat_v, self.flat_p.numel(), lr, b1,
                                      b2, eps, wd, 1.0 / W, self._step)
            return loss
        self.sync_grads()
        if self.flat_p.is_cuda:
            N.check(N.lib().dgb_adam_flat(N.ptr(self.flat_p), N.ptr(self.flat_g), N.ptr(self.flat_m), N.ptr(self.flat_v),
                                          N.ptr(None), N.i64(self.flat_p.numel()), N.f32(lr), N.f32(b1), N.f32(b2),
                                          N.f32(eps), N.f32(wd), N.f32(1.0 / W), N.i32(self._step), N.stream()),
                    "dgb_adam_flat")
            N.count_launch()
        else:
            gr = self.flat_g / W + wd * self.flat_p
            self.flat_m.mul_(b1).add_(gr, alpha=1 - b1)
            self.flat_v.mul_(b2).addcmul_(gr, gr, value=1 - b2)
            c1, c2 = 1 - b1 ** self._step, 1 - b2 ** self._step
            self.flat_p.sub_(lr * (self.flat_m / c1) / (self.flat_v.sqrt() / math.sqrt(c2) + eps))
        return loss

    # -- in-memory snapshot (e.g. around the warm-up steps a CUDA-graph capture needs) ----------------------------------------------
    def snapshot(self):
        """Parameters, moments and step counter as device copies; :meth:`restore` writes them back in place (views stay attached)."""
        snap = {"p": self.flat_p.clone(), "m": self.flat_m.clone(), "v": self.flat_v.clone(), "step": self._step}
        if self.capturable:
            snap["dev_step"] = self._dev_step.clone()
        return snap

    def restore(self, snap) -> None:
        self.flat_p.copy_(snap["p"])
        self.flat_m.copy_(snap["m"])
        self.flat_v.copy_(snap["v"])
        self._step = snap["step"]
        if self.capturable and "dev_step" in snap:
            self._dev_step.copy_(snap["dev_step"])
        _wcache.bump()

    # -- torch.optim.Adam-compatible checkpoint format ---------------------------------------------
    def state_dict(self):
        state = {}
        if self.capturable:  # the device counter is the truth once graphs replay without the host
            self._step = int(self._dev_step.item())
        for i, (p, o, n) in enumerate(self._views):
            state[i] = {"step": torch.tensor(float(self._step)),
                        "exp_avg": self.flat_m[o:o + n].view(p.shape).clone(),
                        "exp_avg_sq": self.flat_v[o:o + n].view(p.shape).clone()}
        groups = [{k: v for k, v in self.param_groups[0].items() if k != "params"}]
        groups[0]["params"] = list(range(len(self._views)))
        return {"state": state, "param_groups": groups}

    def load_state_dict(self, sd):
        for i, (p, o, n) in enumerate(self._views):
            st = sd["state"].get(i, sd["state"].get(str(i)))
            if st is None:
                continue
            self.flat_m[o:o + n].copy_(st["exp_avg"].reshape(-1))
            self.flat_v[o:o + n].copy_(st["exp_avg_sq"].reshape(-1))
            self._step = int(float(st["step"]))
        for k, v in sd["param_groups"][0].items():
            if k != "params":
                self.param_groups[0][k] = v
        _wcache.bump()
        if self.capturable:
            self._dev_step.fill_(self._step)
            self.sync_hyper()
