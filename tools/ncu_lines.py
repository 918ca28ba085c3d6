"""Attribute the per-instruction warp-stall samples of an ncu report to source lines.
usage: python -m tools.ncu_lines <report.ncu-rep> <model image .so> [kernel regex]   (needs ncu, cuobjdump, nvdisasm)
Rows of `ncu --page source` are the SASS instructions of the kernel in order; `nvdisasm -g` of the same cubin gives
the source line of every instruction in the same order."""
import csv
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def sass_lines(so, kernel):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, check=True, capture_output=True)
    cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
    out, cur, on = [], ("?", 0), False
    for l in txt.split("\n"):
        if l.startswith("//--------------------- .text."):
            on = l.split(".text.")[1].split()[0] == kernel
        elif on:
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
            if m:
                if "inlined at" not in l:
                    cur = (os.path.basename(m.group(1)), int(m.group(2)))
            elif re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
                out.append(cur)
    return out


B_SRC = int(os.environ.get("NCU_LINES_SRC", "10"))     # bucket width in source lines: hand-written sources / generated headers
B_GEN = int(os.environ.get("NCU_LINES_GEN", "200"))
TOP = int(os.environ.get("NCU_LINES_TOP", "60"))


def main():
    rep, so = sys.argv[1], sys.argv[2]
    kernel = sys.argv[3] if len(sys.argv) > 3 else "pbgpu_solve_kernel"
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.split("\n")))
    hdr = rows[1]
    data = [r for r in rows[2:] if len(r) == len(hdr)]
    lines = sass_lines(so, kernel)
    assert len(lines) == len(data), (len(lines), len(data))
    sc = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    ex = hdr.index("Instructions Executed")
    agg = defaultdict(lambda: defaultdict(int))
    for (f, ln), r in zip(lines, data):
        key = (f, ln // B_SRC * B_SRC) if f.endswith(".cuh") or f.endswith(".cu") else (f, ln // B_GEN * B_GEN)
        a = agg[key]
        a["instr"] += 1
        a["exec"] += int(r[ex] or 0)
        for k in sc:
            if r[k] not in ("", "0"):
                a[hdr[k][6:]] += int(r[k])
    tot = sum(sum(v for k, v in a.items() if k not in ("instr", "exec")) for a in agg.values())
    print("total samples", tot, "instructions", len(data))
    for key, a in sorted(agg.items(), key=lambda kv: -sum(v for k, v in kv[1].items() if k not in ("instr", "exec")))[:TOP]:
        n = sum(v for k, v in a.items() if k not in ("instr", "exec"))
        top = sorted(((k, v) for k, v in a.items() if k not in ("instr", "exec")), key=lambda kv: -kv[1])[:4]
        print(f"{key[0]:18s} {key[1]:6d}  {100 * n / tot:5.1f}%  instr {a['instr']:6d} exec {a['exec']:10d}  " +
              " ".join(f"{k}={v}" for k, v in top))


if __name__ == "__main__":
    main()
