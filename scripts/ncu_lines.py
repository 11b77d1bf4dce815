#!/usr/bin/env python
"""Join the per-SASS-instruction samples of an ncu report with the source lines of the
kernel (nvdisasm -g line markers), and print the hottest source lines.

  python scripts/ncu_lines.py REPORT.ncu-rep KERNEL_MANGLED_SUBSTRING [--top 50]

Needs the same libmimgpu.so that produced the report (built with -lineinfo)."""
import csv
import io
import re
import subprocess
import sys
import tempfile
from collections import defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "mim_solvers_b200" / "csrc" / "libmimgpu.so"


def sass_lines(kernel_sub):
    """[(file, line)] per instruction of the first kernel whose mangled name contains kernel_sub."""
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", str(LIB)], cwd=d, check=True, capture_output=True)
        out = []
        for cubin in sorted(Path(d).glob("*.cubin")):
            txt = subprocess.run(["nvdisasm", "-g", "-c", str(cubin)], capture_output=True, text=True).stdout
            cur, inside = None, False
            for ln in txt.splitlines():
                if ln.startswith(".text."):
                    if inside and out:
                        return out
                    inside = kernel_sub in ln
                    continue
                if not inside:
                    continue
                m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
                if m:
                    cur = (Path(m.group(1)).name, int(m.group(2)))
                    continue
                if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
                    out.append(cur)
            if out:
                return out
    return out


def main():
    rep, ksub = sys.argv[1], sys.argv[2]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 50
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    h = next(i for i, r in enumerate(rows) if "# Samples" in r)
    hdr, data = rows[h], [r for r in rows[h + 1:] if len(r) > 5]
    iS, iI, iSrc = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
    lines = sass_lines(ksub)
    if len(lines) != len(data):
        print(f"warning: {len(lines)} SASS instructions in the cubin vs {len(data)} in the report", file=sys.stderr)
    agg = defaultdict(lambda: [0, 0])
    tot = 0
    for r, l in zip(data, lines):
        agg[l][0] += int(r[iS])
        agg[l][1] += int(r[iI])
        tot += int(r[iS])
    src_cache = {}
    print(f"total samples {tot}")
    for l, (s, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        text = ""
        if l:
            f = next(iter(ROOT.glob(f"**/{l[0]}")), None)
            if f:
                src_cache.setdefault(f, f.read_text().splitlines())
                text = src_cache[f][l[1] - 1].strip()[:90]
        print(f"{100 * s / max(tot, 1):5.1f}%  inst {n:>12d}  {l}  {text}")


if __name__ == "__main__":
    main()
