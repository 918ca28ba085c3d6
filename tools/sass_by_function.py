"""SASS instruction counts of one kernel of a model image attributed to source functions (through -lineinfo), so that
the code footprint of the hot loop can be read off:  python -m tools.sass_by_function <image.so> [kernel]"""
import bisect
import collections
import os
import re
import subprocess
import sys
import tempfile


def main(path, kernel="pbgpu_solve_kernel"):
    path = os.path.abspath(path)
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", path], cwd=d, capture_output=True)
        cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout.split("\n")
    cur, kern = None, None
    cnt = collections.Counter()
    for ln in txt:
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1), int(m.group(2)))
            continue
        m = re.match(r"\.text\.(\w+):", ln)
        if m:
            kern = m.group(1)
            continue
        if kern == kernel and cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
            cnt[cur] += 1
    funcs = {}
    agg = collections.Counter()
    for (f, line), c in cnt.items():
        if f not in funcs:
            fl = []
            try:
                for i, s in enumerate(open(f).read().split("\n"), 1):
                    m = re.match(r"\s*(?:static |PB_FN |PB_RED_FN |PB_DEV |extern \"C\" |__device__ |__noinline__ |__forceinline__ |__global__ |template <[^>]*> )+[\w\s\*&<>:]*?\b(p\w+)\s*\(", s)
                    if m and not s.rstrip().endswith(";"):
                        fl.append((i, m.group(1)))
            except OSError:
                pass
            funcs[f] = fl
        fl = funcs[f]
        k = bisect.bisect_right([x[0] for x in fl], line) - 1
        agg[(os.path.basename(f), fl[k][1] if k >= 0 else "?")] += c
    tot = sum(agg.values())
    print(f"{kernel}: {tot} instructions, {tot * 16 / 1024:.1f} KB")
    for (f, fn), v in agg.most_common(45):
        print(f"{v:7d} instr {v * 16 / 1024:7.1f} KB  {f}:{fn}")


if __name__ == "__main__":
    main(*sys.argv[1:])
