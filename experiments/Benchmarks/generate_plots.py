"""Mean / std of per-iteration times after 10 warm-up iterations (parity: reference ``generate_plots.py:5-80``)."""
import glob
import json
import sys

import numpy as np


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
