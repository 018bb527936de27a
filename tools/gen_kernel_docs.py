"""Writes ``docs/kernels.md`` from ``include/dgraph_b200.h``: one section per ``/* ---- title ---- */`` block of the header, the
block's comment as prose and every entry point of the block with its signature.  ``python tools/gen_kernel_docs.py [--check]``
(``--check`` exits 1 when the committed page is out of date; ``tests/test_hybrid_and_header.py`` runs it)."""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "dgraph_b200.h")
OUT = os.path.join(ROOT, "docs", "kernels.md")


def _clean(comment: str) -> str:
    lines = [re.sub(r"^\s*\*\s?", "", ln).rstrip() for ln in comment.strip().splitlines()]
    text = " ".join(ln.strip() for ln in lines if ln.strip())
    return re.sub(r"\s*-{4,}\s*", " ", text).strip()


def render() -> str:
    src = open(HDR).read()
    body = src.split('extern "C" {', 1)[1].rsplit("#ifdef __cplusplus", 1)[0]
    out = ["# Kernel library (C ABI)", "",
           "Generated from `include/dgraph_b200.h` by `tools/gen_kernel_docs.py`; the header is the source of truth.  Every entry "
           "point takes raw device pointers and the stream to launch on, returns 0 or a `cudaError_t` value, never synchronises, and "
           "-- unless its description says otherwise -- allocates nothing.  The library is `dgraph_b200/lib/libdgraph_b200.so` "
           "(`python -m dgraph_b200.build`), loadable with `dlopen` / `ctypes` without PyTorch.", ""]
    # tokens: comments and declarations in order
    pos = 0
    pending = ""
    for m in re.finditer(r"/\*(.*?)\*/|((?:const\s+)?[A-Za-z_][\w\s\*]*?\b(dgb_\w+)\s*\([^;]*\);)", body, re.S):
        if m.group(1) is not None:
            c = m.group(1)
            head = re.match(r"\s*-{2,}\s*(.*?)\s*-{2,}", c, re.S)
            if head:  # section header block: title line + optional prose
                title = head.group(1).strip()
                rest = _clean(c[head.end():])
                out += [f"## {title}", ""]
                if rest:
                    out += [rest, ""]
                pending = ""
            else:
                pending = _clean(c)
        else:
            decl = re.sub(r"\s+", " ", m.group(2)).strip()
            out += [f"### `{m.group(3)}`", ""]
            if pending:
                out += [pending, ""]
                pending = ""
            out += ["```c", decl, "```", ""]
    return "\n".join(out).rstrip() + "\n"


def main():
    text = render()
    if "--check" in sys.argv:
        cur = open(OUT).read() if os.path.exists(OUT) else ""
        if cur != text:
            print("docs/kernels.md is out of date: run python tools/gen_kernel_docs.py")
            sys.exit(1)
        return
    with open(OUT, "w") as f:
        f.write(text)
    print(f"wrote {OUT} ({text.count('### ')} entry points)")


if __name__ == "__main__":
    main()
