"""Stacked per-scope epoch-time breakdown from ``{dataset}_timing_report_world{W}.json`` files (parity: reference
``experiments/OGB/plot_timing_reports.py:11-148``).  Prints a table; draws the stacked-bar chart when matplotlib exists."""
import glob
import json
import re
import sys


def plot_stacked_bar_from_dicts(data, x_ticks=None, plot_filename="stacked_bar_plot.png", title="Stacked Bar Plot"):
    """One stacked bar per dictionary of ``data``: keys are timing scopes, values lists of per-epoch milliseconds, segment height
    = mean of the list (scopes ``cache_generation_time`` and ``pre-processing`` are left out, as in the reference).  Returns the
    per-bar means; draws the figure when matplotlib is importable."""
    if not data:
        print("Input data is empty. Cannot generate a plot.")
        return None
    skip = {"cache_generation_time", "pre-processing"}
    keys = [k for k in data[0] if k not in skip]
    means = [{k: (sum(d[k]) / len(d[k]) if d.get(k) else 0.0) for k in keys} for d in data]
    ticks = [str(t) for t in (x_ticks if x_ticks is not None else range(len(data)))]
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        return means
    plt.figure()
    bottom = [0.0] * len(data)
    for k in keys:
        vals = [m[k] for m in means]
        plt.bar(ticks, vals, bottom=bottom, label=k.replace("_", " "))
        bottom = [b + v for b, v in zip(bottom, vals)]
    plt.ylabel("ms / epoch")
    plt.xlabel("ranks")
    plt.title(title)
    plt.legend()
    plt.savefig(plot_filename)
    plt.close()
    return means


def main(log_dir="logs", dataset="products"):
    rows = {}
    for f in sorted(glob.glob(f"{log_dir}/{dataset}_timing_report_world*.json")):
        w = int(re.search(r"world(\d+)", f).group(1))
        t = json.load(open(f))
        rows[w] = {k: (sum(v[1:]) / max(len(v) - 1, 1) if len(v) > 1 else (v[0] if v else 0.0)) for k, v in t.items()}
    if not rows:
        print(f"no timing reports under {log_dir}")
        return
    keys = sorted({k for r in rows.values() for k in r})
    print("world | " + " | ".join(keys))
    for w in sorted(rows):
        print(f"{w:5d} | " + " | ".join(f"{rows[w].get(k, 0.0):8.3f}" for k in keys))
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        return
    bottom = [0.0] * len(rows)
    ws = sorted(rows)
    for k in keys:
        vals = [rows[w].get(k, 0.0) for w in ws]
        plt.bar([str(w) for w in ws], vals, bottom=bottom, label=k)
        bottom = [b + v for b, v in zip(bottom, vals)]
    plt.ylabel("ms / epoch")
    plt.xlabel("ranks")
    plt.title(f"GCN on {dataset}-shaped graph (B200)")
    plt.legend()
    plt.savefig(f"{log_dir}/{dataset}_timing_breakdown.png")


if __name__ == "__main__":
    main(*sys.argv[1:])
