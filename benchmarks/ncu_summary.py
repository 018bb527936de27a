"""Turn an ncu report into the markdown summary committed under profiles/ (B200_PROFILING.md metric names)."""
import csv
import json
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe % of peak"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
        ("launch__registers_per_thread", "regs/thread"), ("lts__t_sector_hit_rate.pct", "L2 hit %"),
        ("smsp__inst_executed.sum", "warp instructions")]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")
    lines = ["| kernel | " + " | ".join(n for _, n in KEYS) + " |", "|---|" + "---|" * len(KEYS)]
    seen = {}
    for r in rows[2:]:
        name = r[ik].split("(")[0][:70]
        vals = []
        for k, _ in KEYS:
            if k in hdr:
                i = hdr.index(k)
                try:
                    vals.append(f"{float(r[i].replace(',', '')):.4g} {units[i]}".strip())
                except ValueError:
                    vals.append(r[i])
            else:
                vals.append("-")
        key = name
        seen[key] = seen.get(key, 0) + 1
        if seen[key] <= 4:
            lines.append(f"| `{name}` | " + " | ".join(vals) + " |")
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
