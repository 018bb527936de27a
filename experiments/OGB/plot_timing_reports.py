"""Stacked per-scope epoch-time breakdown from ``{dataset}_timing_report_world{W}.json`` files (parity: reference
``experiments/OGB/plot_timing_reports.py:11-148``).  Prints a table; draws the stacked-bar chart when matplotlib exists."""
import glob
import json
import re
import sys


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
