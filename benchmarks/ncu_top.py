"""Summarise an ncu `--page source --csv` dump: top stalled SASS instructions and their dominant stall reason."""
import csv
import sys


def main(path, top=30):
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    i_s = hdr.index("# Samples")
    i_src = hdr.index("Source")
    stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    data = []
    for n, r in enumerate(rows[2:]):
        try:
            data.append((int(r[i_s]), n, r))
        except (ValueError, IndexError):
            pass
    tot = sum(d[0] for d in data)
    print(f"total samples {tot}")
    for s, n, r in sorted(data, key=lambda x: -x[0])[:top]:
        st = sorted(((int(r[i] or 0), h) for i, h in stall_cols), reverse=True)[:2]
        print(f"{s:7d} {100.0 * s / max(tot, 1):5.1f}%  line {n:5d}  {r[i_src].strip()[:90]:90s} {st}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30)
