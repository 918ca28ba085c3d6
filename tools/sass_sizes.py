"""Code size of the device functions of a model image (bytes of SASS per function), from the ELF symbol table:
python -m tools.sass_sizes pybamm_b200/_build/pbgpu_model_dfn.so [substring]"""
import re
import subprocess
import sys


def sizes(path):
    txt = subprocess.run(["cuobjdump", "-elf", path], capture_output=True, text=True).stdout
    out, in_sym = {}, False
    for ln in txt.splitlines():
        if ".symtab" in ln:
            in_sym = True
            continue
        f = ln.split()
        if in_sym and len(f) == 7 and f[3] in ("0x2", "0x22", "0x12"):
            try:
                val, size, name = int(f[1], 16), int(f[2], 16), f[-1]
            except ValueError:
                continue
            out[(name, val)] = size
    return out


if __name__ == "__main__":
    s = sizes(sys.argv[1])
    filt = sys.argv[2] if len(sys.argv) > 2 else ""
    rows = sorted(((sz, n, v) for (n, v), sz in s.items() if filt in n), reverse=True)
    for sz, n, v in rows[:40]:
        short = re.sub(r"_ZN\d+_INTERNAL_[0-9a-f]+_\d+_pbgpu_model_cu_[0-9a-f]+\d\d?", " ", n)
        print(f"{sz:9d} B  {sz / 1024:7.1f} KB  @{v:#08x}  {short}")
