"""RGAT trainer (parity: reference ``experiments/OGB-LSC/Trainer.py:47-194``): full-graph epochs, global-sum cross entropy
normalised by the global number of targets, global accuracy via all-reduce, plan-memory report."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

import torch  # noqa: E402

from config import ModelConfig, TrainingConfig  # noqa: E402
from dgraph_b200.data.hetero import RELATIONS  # noqa: E402
from dgraph_b200.models.rgat import CommAwareRGAT  # noqa: E402
from dgraph_b200.nn.dist_bn import GetGlobalVal  # noqa: E402
from dgraph_b200.optim import FusedAdam  # noqa: E402


def print_on_rank_zero(comm, *a):
    if comm.get_rank() == 0:
        print(*a, flush=True)


def print_on_all_ranks(comm, *a):
    """Every rank prints in rank order (barriers in between), reference ``Trainer.py:39-44``."""
    for r in range(comm.get_world_size()):
        if comm.get_rank() == r:
            print(f"[rank {r}]", *a, flush=True)
        comm.barrier()


def stop():
    """Tear the process group down and exit (debugging aid of the reference, ``Trainer.py:25-29``)."""
    import sys

    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
    sys.exit(0)


class Trainer:
    def __init__(self, dataset, comm, device, model_config: ModelConfig = None, training_config: TrainingConfig = None,
                 fused: bool = True):
        self.dataset, self.comm, self.device, self.fused = dataset, comm, device, fused
        self.mc, self.tc = model_config or ModelConfig(), training_config or TrainingConfig()
        self.cdt = torch.bfloat16 if device.type == "cuda" else torch.float32
        self.model = CommAwareRGAT(dataset.num_features, dataset.num_classes, self.mc.hidden_channels, dataset.num_relations,
                                   self.mc.num_layers, self.mc.heads, comm, dropout=self.mc.dropout).to(device)
        # replicated weights + one fused kernel for gradient all-reduce and Adam (the role of DDP + optim.Adam)
        self.optimizer = FusedAdam(self.model.parameters(), lr=self.tc.lr, weight_decay=self.tc.weight_decay, comm=comm,
                                   grad_average=False)
        self.scheduler = torch.optim.lr_scheduler.StepLR(self.optimizer, self.tc.lr_step_size, self.tc.lr_gamma)

    def prepare_data(self):
        self.dataset.prepare_plans(fused=self.fused, reference_layout=not self.fused,
                                   exchange=getattr(self.comm.engine, "exchange_ids", None))
        if not self.fused:
            self.dataset.add_batch_dimension()
        self.dataset.to(self.device)
        plans = self.dataset.patterns if self.fused else self.dataset.comm_plans
        mem = sum(p.memory_usage("MB")["total"] if hasattr(p, "memory_usage") else
                  (p.source_graph_plan.memory_usage("MB")["total"] + p.dest_graph_plan.memory_usage("MB")["total"])
                  for p in plans)
        print_on_rank_zero(self.comm, f"communication plans: {mem:.2f} MB on rank 0")
        return self

    def _forward(self):
        xs = [f.to(self.cdt) for f in self.dataset.features]
        plans = self.dataset.patterns if self.fused else self.dataset.comm_plans
        out = self.model(xs, RELATIONS, plans)
        return out[0] if out.dim() == 3 else out

    def train(self, epochs=None):
        epochs = epochs or self.tc.epochs
        train_idx = self.dataset.get_mask("train").to(self.device)
        target = self.dataset.get_target("train")
        n_global = GetGlobalVal(train_idx.numel(), self.comm, self.device)
        times = []
        for ep in range(epochs):
            self.model.train()
            self.comm.barrier()
            t0 = time.perf_counter()
            self.optimizer.zero_grad()
            out = self._forward()
            loss = torch.nn.functional.cross_entropy(out.index_select(0, train_idx).float(), target, reduction="sum") / n_global
            loss.backward()
            self.optimizer.step()
            self.scheduler.step()
            if self.device.type == "cuda":
                torch.cuda.synchronize()
            self.comm.barrier()
            times.append((time.perf_counter() - t0) * 1e3)
            gl = GetGlobalVal(float(loss.detach()), self.comm, self.device)
            if ep % 10 == 0 or ep == epochs - 1:
                print_on_rank_zero(self.comm, f"epoch {ep:4d} | loss {gl:.4f} | {times[-1]:.1f} ms")
        return times

    @torch.no_grad()
    def evaluate(self, split="valid"):
        self.model.eval()
        idx = self.dataset.get_mask(split).to(self.device)
        target = self.dataset.get_target(split)
        out = self._forward().index_select(0, idx)
        correct = (out.argmax(1) == target).sum().item()
        total = GetGlobalVal(idx.numel(), self.comm, self.device)
        return GetGlobalVal(correct, self.comm, self.device) / max(total, 1.0)
