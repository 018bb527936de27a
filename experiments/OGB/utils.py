"""Rank-0 experiment logs and small metrics helpers (parity: reference ``experiments/OGB/utils.py:7-64``)."""
import os


def make_experiment_log(fname: str, rank: int):
    if rank == 0:
        os.makedirs(os.path.dirname(fname) or ".", exist_ok=True)
        open(fname, "w").close()


def write_experiment_log(log: str, fname: str, rank: int):
    if rank == 0:
        with open(fname, "a") as f:
            f.write(log + "\n")


def dist_print_ephemeral(msg: str, rank: int):
    if rank == 0:
        print(msg, end="\r", flush=True)


def calculate_accuracy(pred, labels) -> float:
    if labels.numel() == 0:
        return 0.0
    return float((pred.argmax(1) == labels.reshape(-1)).float().mean() * 100.0)


def safe_plot(x, ys, labels, title, path):
    """Mean +- std trajectory plot; silently skipped when matplotlib is missing (it is on this image)."""
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        import numpy as np
    except ImportError:
        return False
    plt.figure()
    for y, l in zip(ys, labels):
        y = np.asarray(y)
        m, s = y.mean(0), y.std(0)
        plt.plot(x, m, label=l)
        plt.fill_between(x, m - s, m + s, alpha=0.3)
    plt.title(title)
    plt.legend()
    plt.savefig(path)
    plt.close()
    return True


def cleanup():
    """Tear down the default process group if there is one (reference ``utils.py:7-9``)."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


def safe_create_dir(directory, rank):
    if rank == 0:
        os.makedirs(directory, exist_ok=True)


def visualize_trajectories(trajectories, title, figsave, rank):
    """Rank 0 plots mean +- std over runs of a per-epoch metric (``[runs, epochs]``) and saves the figure."""
    if rank != 0:
        return False
    import numpy as np

    t = np.asarray(trajectories)
    return safe_plot(np.arange(t.shape[-1]), [np.atleast_2d(t)], [title], title, figsave)
