"""Source generation and compilation drivers (host side, run once per model).

Plays the role of ``pybamm.codegen.compilation.aot_compile`` in the reference
(packages/pybamm/src/pybamm/codegen/compilation.py:61-196: CasADi CodeGenerator -> C ->
gcc -> shared object, cached on disk), but the product target is an sm_100a kernel image:

  bundle (.npz) --codegen--> <name>_cuda.h --nvcc--> pybamm_b200/_build/pbgpu_model_<name>.so
                \\-codegen--> <name>_c.h    --gcc---> oracle/_build/oracle_<name>.so   (tests / CPU baseline only)
"""
from __future__ import annotations

import hashlib
import json
import os
import subprocess

from .bundle import ModelBundle
from . import codegen

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
BUILD = os.path.join(PKG, "_build")
CSRC = os.path.join(PKG, "csrc")
MODELS = os.path.join(PKG, "models")
ORACLE = os.path.join(ROOT, "oracle")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--cudart", "shared",
              "-Xcompiler", "-fPIC", "--shared", "-Xptxas", "-v"]


def load_bundle(name_or_path) -> ModelBundle:
    path = name_or_path
    if not os.path.exists(path):
        path = os.path.join(MODELS, f"{name_or_path}.npz")
    return ModelBundle.load(path)


def _write_if_changed(path, text):
    if os.path.exists(path):
        with open(path) as f:
            if f.read() == text:
                return False
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        f.write(text)
    return True


def _src_digest(paths):
    h = hashlib.sha256()
    for p in paths:
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def generate(bundle: ModelBundle, dialect: str, outdir=BUILD, tag="", extra_samples=None):
    text, info = codegen.generate(bundle, dialect, extra_samples=extra_samples)
    path = os.path.join(outdir, f"{bundle.name}{tag}_{dialect}.h")
    _write_if_changed(path, text)
    with open(os.path.join(outdir, f"{bundle.name}{tag}_{dialect}.json"), "w") as f:
        json.dump(info, f, indent=1)
    return path, info


def build_oracle(bundle: ModelBundle, force=False, fast=False):
    """Compile the CPU oracle for a bundle (tests / bench cpu_baseline only).
    fast: the CPU-baseline build (-O3 -march=native: compiled on the machine that runs it, never shipped)."""
    outdir = os.path.join(ORACLE, "_build")
    os.makedirs(outdir, exist_ok=True)
    header, _ = generate(bundle, "c", outdir)
    out = os.path.join(outdir, f"oracle_{bundle.name}{'_fast' if fast else ''}.so")
    srcs = [os.path.join(ORACLE, f) for f in ("ida_oracle.c", "model_shim.c", "pbo_model.h", "Makefile")] + [header]
    stamp = out + ".stamp"
    dig = _src_digest(srcs)
    if fast:
        import platform

        cpu = ""
        try:
            cpu = next(ln for ln in open("/proc/cpuinfo") if ln.startswith("model name")).strip()
        except Exception:
            pass
        dig += hashlib.sha256((platform.machine() + cpu).encode()).hexdigest()[:8]
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == dig:
        return out
    cmd = ["make", "-s", "-C", ORACLE, "-B", f"MODEL={bundle.name}", f"HEADER={header}", f"OUT={out}"] + (["FAST=1"] if fast else [])
    subprocess.run(cmd, check=True)
    with open(stamp, "w") as f:
        f.write(dig)
    return out


def lu_order(bundle: ModelBundle):
    """(perm_r, perm_c) of the static sparse LU of a bundle: read from the code generator's record of the last build of
    the model image (<name>_cuda.json), recomputed when there is none."""
    path = os.path.join(BUILD, f"{bundle.name}_cuda.json")
    try:
        lu = json.load(open(path))["lu"]
        return lu["perm_r"], lu["perm_c"]
    except Exception:
        from . import symbolic_lu as slu

        sym = slu.analyse(bundle.n_states, bundle.jac_rows, bundle.jac_cols, codegen.sample_matrices(bundle),
                          team=codegen.default_team(bundle.n_states))
        return [int(x) for x in sym.perm_r], [int(x) for x in sym.perm_c]


def build_model(bundle: ModelBundle, force=False, team=None, verbose=False, profile=False, tag="", defines=(), extra_flags=(),
                gvec=None, min_teams=None, extra_samples=None):
    """Generate + nvcc-compile the sm_100a solver image of a bundle. Returns the .so path.
    extra_samples: see codegen.generate (a re-pivoted image; give it a tag of its own)."""
    os.makedirs(BUILD, exist_ok=True)
    header, info = generate(bundle, "cuda", BUILD, tag=tag, extra_samples=extra_samples)
    out = os.path.join(BUILD, f"pbgpu_model_{bundle.name}{('_prof' + (str(int(profile)) if int(profile) > 1 else '')) if profile else ''}{tag}.so")
    kernel_srcs = [os.path.join(CSRC, f) for f in ("pbgpu_model.cu", "pbgpu_ida.cuh", "pbgpu_team.cuh")]
    inc = os.path.join(ROOT, "include")
    srcs = kernel_srcs + [header, os.path.join(inc, "pbgpu.h")]
    if team is None:
        team = default_team(bundle)
    if gvec is None:
        # streamed vectors in global scratch for large models; 2: residual vector and solve stream as well (pbgpu_ida.cuh)
        gvec = 0 if bundle.n_states <= 256 else (1 if bundle.n_states <= 2048 else 2)
    if min_teams is None:
        # register cap.  Large models: two 256-thread teams per SM (128 registers).  Tiny one-warp models (SPM, 42 states):
        # without a cap the kernel takes 255 registers and only 8 warps fit an SM; 16 warps at 128 registers measured +25 %
        # (24 / 32 warps spill too much).  In between (SPMe, 102 states) shared memory binds first and the cap only costs.
        # Which of the two wins depends on the batch size, so such an image carries both kernels (PB_MIN_TEAMS_WIDE) and the
        # launch picks: uncapped while every system of the batch is resident at once, capped beyond that.
        min_teams = 2 if gvec == 1 else 1
    wide = 16 if (not gvec and team == 32 and bundle.n_states <= 64 and min_teams == 1) else 0
    if team == 32 and bundle.n_states > 64 and not defines and os.environ.get("PBGPU_AUTO_DEFINES", "1") != "0":
        # One-warp models whose vectors are three or more elements per lane (SPMe: 102 states): nine warps per SM run nine
        # different places of ~150 KB of per-step code against a 32 KB instruction cache (no_instruction is half of all stall
        # samples), so code size is what counts.  Rolled vector loops and out-of-line reductions: 23.8k -> 26.2k sims/s.  With
        # two elements per lane (SPM) the unrolled form stays faster (161k against 148k).
        defines = ("PB_VEC_ROLLED", "PB_TEAM_INLINE_RED=0")
    dig = _src_digest(srcs) + f"-t{team}-g{gvec}-m{min_teams}-w{wide}-p{int(profile)}-{'-'.join(defines)}-{'-'.join(extra_flags)}-" + hashlib.sha256(" ".join(NVCC_FLAGS).encode()).hexdigest()[:8]
    stamp = out + ".stamp"
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == dig:
        return out
    cmd = [NVCC] + NVCC_FLAGS + [f"-DPBM_HEADER=\"{header}\"", f"-DPB_TEAM_SIZE={team}", f"-DPB_GLOBAL_VEC={gvec}", f"-DPB_MIN_TEAMS={min_teams}"] + ([f"-DPB_MIN_TEAMS_WIDE={wide}"] if wide else []) + [f"-I{inc}", f"-I{CSRC}",
                                 kernel_srcs[0], "-o", out] + ([f"-DPB_PROFILE={int(profile)}"] if profile else []) + [f"-D{d}" for d in defines] + list(extra_flags)
    r = subprocess.run(cmd, capture_output=True, text=True)
    log = r.stdout + r.stderr
    with open(out[:-3] + ".ptxas.log", "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if verbose or r.returncode != 0:
        print(log)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for model {bundle.name}")
    with open(stamp, "w") as f:
        f.write(dig)
    return out


def default_team(bundle: ModelBundle) -> int:
    """Threads per system (codegen.default_team: the SEQ solve stream is laid out for it)."""
    return codegen.default_team(bundle.n_states)


def build_api(force=False):
    """Compile the generic C-ABI host library libpbgpu.so (include/pbgpu.h)."""
    os.makedirs(BUILD, exist_ok=True)
    out = os.path.join(BUILD, "libpbgpu.so")
    src = os.path.join(CSRC, "pbgpu_api.cpp")
    inc = os.path.join(ROOT, "include")
    dig = _src_digest([src, os.path.join(inc, "pbgpu.h")])
    stamp = out + ".stamp"
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == dig:
        return out
    cuda_home = os.path.dirname(os.path.dirname(NVCC))
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", f"-I{inc}", f"-I{cuda_home}/include", src, "-o", out,
           f"-L{cuda_home}/lib64", "-lcudart", "-ldl", f"-Wl,-rpath,{cuda_home}/lib64"]
    subprocess.run(cmd, check=True)
    with open(stamp, "w") as f:
        f.write(dig)
    return out


def build_peak(force=False):
    """Compile the FP64 peak microbenchmark used by bench.py's roofline (include/pbgpu_peak.h)."""
    os.makedirs(BUILD, exist_ok=True)
    out = os.path.join(BUILD, "libpbgpu_peak.so")
    src = os.path.join(CSRC, "pbgpu_peak.cu")
    inc = os.path.join(ROOT, "include")
    dig = _src_digest([src, os.path.join(inc, "pbgpu_peak.h")])
    stamp = out + ".stamp"
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == dig:
        return out
    cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "--cudart", "shared", "-Xcompiler", "-fPIC",
           "--shared", f"-I{inc}", src, "-o", out]
    subprocess.run(cmd, check=True)
    with open(stamp, "w") as f:
        f.write(dig)
    return out
