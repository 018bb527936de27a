"""Mean / std of per-iteration times after 10 warm-up iterations (parity: reference ``generate_plots.py:5-80``)."""
import glob
import json
import sys

import numpy as np


def get_stats(_file, warmup=10):
    """``(mean, std)`` of the per-iteration times in a ``.npy`` file after dropping the first ``warmup`` iterations."""
    t = np.load(_file)[warmup:]
    return float(np.mean(t)), float(np.std(t))


def _bar(labels, means, stds, title, path):
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        return False
    plt.figure()
    plt.bar(labels, means, yerr=stds)
    plt.ylabel("ms / call")
    plt.title(title)
    plt.savefig(path)
    plt.close()
    return True


def generate_plots(backend: str, log_dir: str = "logs", engine: str = "*"):
    """Gather / scatter mean +- std of one backend from the files ``TestNCCL.py`` / ``TestNVSHMEM.py`` wrote (rank 0)."""
    res = {}
    for op in ("gather", "scatter"):
        files = sorted(glob.glob(f"{log_dir}/legacy_{op}_cache_{engine}_rank0.npy"))
        if files:
            res[op] = get_stats(files[0])
    if res:
        _bar(list(res), [v[0] for v in res.values()], [v[1] for v in res.values()], f"{backend.upper()} gather / scatter",
             f"{log_dir}/{backend}_comm_times.png")
    return res


def generate_cache_comparison_plot(log_dir: str = "logs", engine: str = "*"):
    """With-cache vs without-cache times of the legacy API (``TestNCCL.py --benchmark_cache``)."""
    res = {}
    for op in ("gather", "scatter"):
        for tag in ("cache", "nocache"):
            files = sorted(glob.glob(f"{log_dir}/legacy_{op}_{tag}_{engine}_rank0.npy"))
            if files:
                res[f"{op} ({tag})"] = get_stats(files[0])
    if res:
        _bar(list(res), [v[0] for v in res.values()], [v[1] for v in res.values()], "plan cache on / off",
             f"{log_dir}/cache_comparison.png")
    return res


def main(log_dir="logs"):
    out = {}
    for f in sorted(glob.glob(f"{log_dir}/*_times_*.npy")):
        t = np.load(f)[10:]
        out[f] = {"mean_ms": float(t.mean()), "std_ms": float(t.std()), "n": int(t.size)}
    print(json.dumps(out, indent=1))
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt

        plt.bar(range(len(out)), [v["mean_ms"] for v in out.values()], yerr=[v["std_ms"] for v in out.values()])
        plt.savefig(f"{log_dir}/comm_times.png")
    except ImportError:
        pass


if __name__ == "__main__":
    main(*sys.argv[1:])
