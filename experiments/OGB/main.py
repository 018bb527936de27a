"""Full-graph GCN training on OGB node-property datasets (parity: reference ``experiments/OGB/main.py:50-362``).

``torchrun --nnodes 1 --nproc-per-node N experiments/OGB/main.py --backend nccl --dataset products --epochs 10``

Same flags and output files as the reference (``{log_dir}/{dataset}_world{W}_run{r}_{training_loss,validation_loss,
validation_accuracy,test_results,training_times,runtime_experiment}.log``, ``{dataset}_timing_report_world{W}.json``,
trajectory PNGs when matplotlib exists).  Adds ``--scale`` (synthetic stand-in size), ``--save`` / ``--resume``
(the reference's Readme promises best-model saving but never implements it).
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _common import cleanup, init_distributed  # noqa: E402

import numpy as np  # noqa: E402
import torch  # noqa: E402

from dgraph_b200.ops.loss import masked_cross_entropy  # noqa: E402
from dgraph_b200.utils.cli import run_cli  # noqa: E402
from dgraph_b200.utils.tracing import KernelTrace  # noqa: E402
from dgraph_b200.utils.watchdog import StepWatchdog, health_check, inject_fault  # noqa: E402
from utils import calculate_accuracy, dist_print_ephemeral, make_experiment_log, safe_plot, write_experiment_log  # noqa: E402


_OPTS = {"watchdog_s": 0.0, "health_every": 0, "trace": ""}


def _run_experiment(dataset, comm, lr, epochs, log_prefix, hidden_dims, num_classes, device, rank, save="", resume=""):
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.parallel import HaloExchange
    from dgraph_b200.utils import TimingReport
    from dgraph_b200.utils.checkpoint import load_checkpoint, save_checkpoint

    x, y, pat = dataset[0]
    cdt = torch.bfloat16 if device.type == "cuda" else torch.float32
    in_dim = x.shape[1]
    pad = (-in_dim) % 8
    y = y.to(device)
    pat.to(device)
    exchanger = HaloExchange(comm)
    # the features live where the first layer's exchange can use them in place (symmetric memory on the peer engine)
    xs = exchanger.symmetric_rows(pat, in_dim + pad, cdt, device=device)
    xs.copy_(torch.nn.functional.pad(x, (0, pad)).to(device, cdt))
    x = xs
    masks = {k: v.to(device) for k, v in dataset.get_masks().items()}
    model = CommAwareGCN(in_dim, hidden_dims, num_classes, exchanger, comm, compute_dtype=cdt).to(device)
    opt = FusedAdam(model.parameters(), lr=lr, comm=comm)
    crit = torch.nn.CrossEntropyLoss()
    train_idx = torch.nonzero(masks["train_mask"], as_tuple=True)[0]  # index form once, outside the epoch loop
    start = load_checkpoint(resume, model, opt, map_location=device) if resume else 0
    tr_loss, va_loss, va_acc, times = [], [], [], []
    for n in ("training_loss", "validation_loss", "validation_accuracy"):
        make_experiment_log(f"{log_prefix}_{n}.log", rank)
    best = -1.0
    wd = StepWatchdog(_OPTS["watchdog_s"]).start() if _OPTS["watchdog_s"] > 0 else None
    for epoch in range(start, start + epochs):
        model.train()
        if wd is not None:
            wd.beat()
        trace = KernelTrace(cuda_events=True) if (_OPTS["trace"] and epoch == start + 1 and rank == 0) else None
        if trace is not None:
            trace.__enter__()
        comm.barrier()
        if device.type == "cuda":
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        opt.zero_grad()
        with TimingReport("forward"):
            out = model(x, pat)
        loss = masked_cross_entropy(out, y, train_idx)  # == crit(out[train_mask], y[train_mask]), fused, no host sync
        with TimingReport("backward"):
            loss.backward()
        opt.step()
        comm.barrier()
        if device.type == "cuda":
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1))
        if trace is not None:
            trace.__exit__(None, None, None)
            trace.to_chrome_trace(_OPTS["trace"])
            print("\n" + "\n".join(f"{k:40s} x{v['launches']:3d} {v['gpu_ms']:8.3f} ms" for k, v in trace.summary().items()))
        loss = inject_fault(rank, epoch, loss)
        if _OPTS["health_every"] > 0 and (epoch + 1) % _OPTS["health_every"] == 0 and not health_check(comm, loss):
            if save:
                save_checkpoint(save + ".emergency", model, opt, step=epoch, rank=rank)
            raise RuntimeError(f"health check failed at epoch {epoch} (non-finite loss or barrier timeout on some rank)")
        tr_loss.append(float(loss.detach()))
        dist_print_ephemeral(f"Epoch {epoch:4d} | loss: {tr_loss[-1]:.4f} | {times[-1] if times else 0:.1f} ms", rank)
        write_experiment_log(str(tr_loss[-1]), f"{log_prefix}_training_loss.log", rank)
        model.eval()
        with torch.no_grad():
            vo = out[masks["val_mask"]].float()
            vl = y[masks["val_mask"]]
            va_loss.append(float(crit(vo, vl.reshape(-1))) if vl.numel() else 0.0)
            va_acc.append(calculate_accuracy(torch.log_softmax(vo, 1), vl))
        write_experiment_log(str(va_loss[-1]), f"{log_prefix}_validation_loss.log", rank)
        write_experiment_log(f"Validation Accuracy: {va_acc[-1]:.2f}", f"{log_prefix}_validation_accuracy.log", rank)
        if save and va_acc[-1] > best:
            best = va_acc[-1]
            save_checkpoint(save, model, opt, step=epoch + 1, rank=rank)
    if wd is not None:
        wd.stop()
    model.eval()
    with torch.no_grad():
        to = model(x, pat)[masks["test_mask"]].float()
        tl = y[masks["test_mask"]]
        test_loss = float(crit(to, tl.reshape(-1))) if tl.numel() else 0.0
        test_acc = calculate_accuracy(torch.log_softmax(to, 1), tl)
    make_experiment_log(f"{log_prefix}_test_results.log", rank)
    write_experiment_log("loss,accuracy", f"{log_prefix}_test_results.log", rank)
    write_experiment_log(f"{test_loss},{test_acc}", f"{log_prefix}_test_results.log", rank)
    make_experiment_log(f"{log_prefix}_training_times.log", rank)
    for t in times:
        write_experiment_log(str(t), f"{log_prefix}_training_times.log", rank)
    avg = float(np.mean(times[1:])) if len(times) > 1 else (times[0] if times else 0.0)
    make_experiment_log(f"{log_prefix}_runtime_experiment.log", rank)
    write_experiment_log(f"Average time per epoch (excl. first): {avg:.4f} ms", f"{log_prefix}_runtime_experiment.log", rank)
    if rank == 0:
        print(f"\nTest  | loss: {test_loss:.4f} | accuracy: {test_acc:.2f}% | avg epoch {avg:.2f} ms")
    return np.array(tr_loss), np.array(va_loss), np.array(va_acc)


def main(backend: str = "nccl", dataset: str = "arxiv", epochs: int = 10, lr: float = 0.001, runs: int = 1,
         hidden_dims: int = 256, log_dir: str = "logs", node_rank_placement_file: str = "", root_dir: str = "data",
         scale: float = 1.0, save: str = "", resume: str = "", timing_report: bool = True, watchdog_s: float = 0.0,
         health_every: int = 0, trace: str = "", partition: str = "contiguous"):
    """Reference flags (``experiments/OGB/main.py:235-362``) plus: ``scale`` (synthetic graph size), ``save`` / ``resume``
    (best-validation checkpoint), ``watchdog_s`` (abort the rank after that many seconds without an epoch), ``health_every``
    (all-rank NaN / timeout check every N epochs), ``trace`` (write a chrome trace of epoch 1's native kernel launches)."""
    _OPTS.update(watchdog_s=float(watchdog_s), health_every=int(health_every), trace=trace)
    from dgraph_b200.data.ogbn_datasets import DGraphOGBDataset
    from dgraph_b200.utils import TimingReport

    assert dataset in ("arxiv", "products"), "dataset must be arxiv or products"
    comm, rank, world, dev = init_distributed(backend)
    TimingReport.init(comm, enabled=timing_report)
    placement = torch.load(node_rank_placement_file, weights_only=False) if node_rank_placement_file else None
    ds = DGraphOGBDataset(dataset, comm, placement, root_dir, scale=scale, device=dev, partition=partition)
    if rank == 0:
        print(f"{dataset}{' (synthetic stand-in)' if ds.synthetic else ''}: V={ds.num_nodes} E={ds.num_edges} world={world}")
    os.makedirs(log_dir, exist_ok=True)
    curves = []
    for r in range(runs):
        curves.append(_run_experiment(ds, comm, lr, epochs, f"{log_dir}/{dataset}_world{world}_run{r}", hidden_dims,
                                      ds.num_classes, dev, rank, save, resume))
    if rank == 0:
        with open(f"{log_dir}/{dataset}_timing_report_world{world}.json", "w") as f:
            json.dump(TimingReport.get_timers(), f)
        x = np.arange(epochs)
        for i, name in enumerate(("training_loss", "validation_loss", "validation_accuracy")):
            safe_plot(x, [np.stack([c[i] for c in curves])], [name], f"{dataset} {name}", f"{log_dir}/{dataset}_{name}_world{world}.png")
    cleanup(comm)


if __name__ == "__main__":
    run_cli(main)
