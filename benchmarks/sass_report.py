"""SASS evidence per kernel, from the built library (no GPU needed): `python benchmarks/sass_report.py > profiles/sass_report.md`.

For every `__global__` entry of `dgraph_b200/lib/libdgraph_b200.so`: instruction count, register / spill numbers from the ptxas
logs, and the mnemonics that prove which hardware path it uses (B200_PROFILING.md: UTCHMMA = tcgen05.mma, UTMALDG / UTMASTG = TMA,
LDTM = tcgen05.ld, SYNCS = mbarrier, HADD2/HSET2/HMNMX2.BF16_V2 = packed bf16 math, *.STRONG.SYS = system-scope flags)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dgraph_b200", "lib", "libdgraph_b200.so")
PROOF = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "SYNCS", "HADD2.BF16_V2", "HFMA2.BF16_V2", "HFMA2.BF16_V2.RELU", "FHADD.BF16", "HSET2.BF16_V2",
         "HMNMX2.BF16_V2", "LDG.E.NA.128", "LDG.E.128", "STG.E.EF", "RED.E", "ATOMG", "STRONG.SYS", "R2P", "SHFL", "MUFU"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def ptxas_info():
    info = {}
    for f in sorted(os.listdir(os.path.join(ROOT, "build", "obj"))):
        if not f.endswith(".ptxas.log"):
            continue
        cur = None
        for ln in open(os.path.join(ROOT, "build", "obj", f)):
            m = re.search(r"Compiling entry function '(\S+)'", ln)
            if m:
                cur = m.group(1)
                info[cur] = {"file": f.replace(".ptxas.log", ".cu")}
            m = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", ln)
            if m and cur:
                info[cur]["spill"] = f"{m.group(1)}/{m.group(2)}"
            m = re.search(r"Used (\d+) registers", ln)
            if m and cur:
                info[cur]["regs"] = m.group(1)
    return info


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for ln in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m and cur:
            kernels[cur][m.group(1)] += 1
    names = demangle(list(kernels))
    info = ptxas_info()
    print("| kernel | source | SASS instructions | registers | spill st/ld (B) | proof mnemonics (count) |")
    print("|---|---|---|---|---|---|")
    for k, c in sorted(kernels.items(), key=lambda kv: (info.get(kv[0], {}).get("file", ""), names[kv[0]])):
        total = sum(c.values())
        proof = []
        for p in PROOF:
            n = sum(v for op, v in c.items() if p in op)
            if n:
                proof.append(f"{p} x{n}")
        nm = re.sub(r"\(.*", "", names[k]).replace("void ", "").replace("dgb::", "")
        i = info.get(k, {})
        print(f"| `{nm[:80]}` | {i.get('file', '?')} | {total} | {i.get('regs', '?')} | {i.get('spill', '?')} | {', '.join(proof)} |")


if __name__ == "__main__":
    sys.exit(main())
