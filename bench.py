#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric: DFN simulations / second (10k input sets, rtol = atol = 1e-8).

  python bench.py --gpus N --steps K --warmup W                (N > 1: launched by torch.distributed.run, one rank per GPU)
  python bench.py --impl reference --gpus N --steps K --warmup W     (CPU arm: the reference algorithm on the host cores)
  python bench.py --scaling strong --gpus N ...                (the metric's own batch: 10 000 sets in TOTAL, split over the ranks)
  python bench.py --inproc-devices N                           (one process, one BatchedIDASolver(devices=range(N)).solve call)
  python bench.py --model spm --n-sys 1000 | --model spme      (BASELINE configs 1 and 2)

A "step" is one batched solve of the workload: DFN (Chen2020, default mesh, 962 states), constant-current discharge,
t_eval = [0, 3700] s, t_interp = linspace(0, 3600, 361), 10 000 input sets per GPU varying four parameters (current, both
particle diffusivities, initial electrolyte concentration), output_variables = ["Voltage [V]"].
The batch is sharded by rank with no collective inside the integration (weak scaling: every rank integrates its own 10k
sets; --scaling strong: shard.partition of one 10k batch).

`value`   : device-resident inputs, pbgpu_solve_device on a CUDA stream, CUDA events around each step.
`e2e`     : the public call BatchedIDASolver.solve(t_eval, inputs=[{...}, ...]) with host buffers: flattening the input
            dicts, H2D, kernel, D2H and building the Solution objects are all inside the timed region.
`roofline`: FP64 pipe. achieved = W(algorithmic flop per simulation) * sims / kernel time; peak = DFMA rate
            measured live on this box by pybamm_b200/csrc/pbgpu_peak.cu.  The HBM fraction is reported beside it.
`cpu_baseline`: the CPU restatement of the reference algorithm (oracle/ida_oracle.c, "port": the reference IDAKLU cannot
            be built here) in its baseline build -- static-pattern sparse LU (what KLU's refactor does), -O3 -march=native,
            OpenMP over systems, all host cores -- on a bounded sample of the same input sets; its IDA counters feed W.
`parity`  : the first --parity-sample systems against the checker build of the oracle (banded LU with partial pivoting) at
            1e-8 and against the oracle at 1e-11 ("truth").  The line's `parity_ok` is False -- and the exit code 3 -- when a
            bar is exceeded (BASELINE.md section 6 states the bars).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

T_EVAL = [0.0, 3700.0]
T_INTERP = np.linspace(0.0, 3600.0, 361)
RTOL = ATOL = 1e-8
UNIT = "simulations/s"
V_BAR, T_BAR = 1e-6, 1e-3          # north_star: terminal voltage within 1e-6 V, event times within 1e-3 s


def metric_name(args):
    if args.model == "dfn":
        return "DFN simulations/sec (10k input sets)"
    if args.model == "dfn_thermal40":
        return f"DFN + lumped thermal (40-point meshes) simulations/sec ({args.n_sys} input sets)"
    return f"{args.model.upper()} simulations/sec ({args.n_sys} input sets)"


def sample_inputs(bundle, n, rng):
    """deterministic input sets (numpy default_rng)."""
    cols = []
    for name in bundle.input_names:
        if name == "Current function [A]":
            cols.append(rng.uniform(2.5, 7.5, n))
        elif name == "Negative particle diffusivity [m2.s-1]":
            cols.append(10 ** rng.uniform(-14, -13, n))
        elif name == "Positive particle diffusivity [m2.s-1]":
            cols.append(10 ** rng.uniform(-15, -14, n))
        elif name == "Initial concentration in electrolyte [mol.m-3]":
            cols.append(rng.uniform(900, 1100, n))
        else:
            raise KeyError(name)
    return np.ascontiguousarray(np.column_stack(cols))


def workload_config(args, world, n_total):
    if args.model == "dfn_thermal40":
        what = (f"DFN + lumped thermal, Chen2020, x_n = x_s = x_p = r_n = r_p = 40 points (3523 states), constant-current discharge, "
                f"{n_total} input sets sweeping the current over 0.5C-1.5C")
    else:
        what = (f"{args.model.upper()} Chen2020 default mesh, 1C-range constant-current discharge, "
                f"{n_total} input sets varying {4 if args.model == 'dfn' else 2} parameters"
                + (", 3.0 V cut-off" if args.model == "spme" else ""))
    return {
        "workload": what,
        "model_states": None,  # filled by the caller
        "n_sys_per_gpu": args.n_sys if args.scaling == "weak" else None,
        "n_sys_total": n_total,
        "rtol": RTOL, "atol": ATOL,
        "t_eval": T_EVAL, "t_interp": "linspace(0, 3600, 361)",
        "output": "Voltage [V]",
        "sharding": f"batch rows over {world} rank(s), no collective",
        "l2": "flushed between timed steps (256 MiB write)",
        "launch_order": "rows handed to the teams by descending predicted step count (log-linear fit on a previous solve), "
                        "results in row order",
    }


def reference_install_probe():
    """is a real reference install importable on this box?  (It replaces the port if it ever is.)"""
    code = ("import sys; sys.path.insert(0, 'baseline/_ref')\n"
            "import importlib\n"
            "out = {}\n"
            "for m in ('pybamm', 'pybammsolvers', 'casadi'):\n"
            "    try:\n"
            "        importlib.import_module(m); out[m] = True\n"
            "    except Exception as e:\n"
            "        out[m] = type(e).__name__\n"
            "print(out)\n")
    try:
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120, cwd=ROOT)
        return {"importable": "True" in r.stdout and "Error" not in r.stdout, "detail": r.stdout.strip()[:200]}
    except Exception as e:
        return {"importable": False, "detail": str(e)[:200]}


def ncu_dram_traffic(n_sys):
    """dram__bytes_read.sum + dram__bytes_write.sum of pbgpu_solve_kernel from the newest committed `ncu --set full`
    capture (profiles/r0X_dfn_full_ncu_raw.csv: `ncu --page raw --csv` of one launch with one DFN system per team, so the
    grid size is the number of systems), scaled to this launch's system count.  Returns (bytes per launch or None, provenance)."""
    import csv
    import glob

    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r0*_dfn_full_ncu_raw.csv")), reverse=True):
        try:
            rows = [r for r in csv.reader(open(path)) if r]
            hdr, units, vals = rows[0], rows[1], rows[2]
            col = {h: k for k, h in enumerate(hdr)}
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            total = sum(float(vals[col[m]].replace(",", "")) * scale[units[col[m]]]
                        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            n_cap = int(float(vals[col["launch__grid_size"]].replace(",", "")))
            per_sim = total / n_cap
            return per_sim * n_sys, (f"{os.path.relpath(path, ROOT)}: ncu --set full, one launch of {n_cap} systems "
                                     f"({per_sim / 1e3:.1f} KB per simulation), scaled to {n_sys} systems")
        except Exception:
            continue
    return None, "no ncu capture found"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax.append(float(f[1]))
                power.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "power_w_max": float(max(power)),
                "samples": len(sm), "reasons": sorted(reasons)}


def measure_fp64_peak(device):
    from pybamm_b200 import build

    lib = C.CDLL(build.build_peak())
    lib.pbgpu_measure_fp64_peak.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_float)]
    tf, ms = C.c_double(0.0), C.c_float(0.0)
    if lib.pbgpu_measure_fp64_peak(device, 4096, C.byref(tf), C.byref(ms)):
        raise RuntimeError("FP64 peak microbenchmark failed")
    return float(tf.value)


def work_model(desc, stats_mean, n_out, n_states):
    """Algorithmic FP64 work per simulation: static costs of the generated functions and
    of the static sparse LU times the IDA counters of the reference algorithm (CPU oracle) for the same inputs."""
    nst, nre, nsetups, netf, nni, ncfn, nje, nge = [float(x) for x in stats_mean]
    f_res, f_jac, f_fac, f_slv = (float(desc.flops[k]) for k in range(4))
    kbar = 3.5   # mean BDF order over a discharge (orders 2..5)
    f_vec = (4 * kbar + 22) * n_states
    f_dky = 2 * (kbar + 1) * n_states
    n_att = nst + netf + ncfn
    w = nre * f_res + nje * f_jac + nsetups * f_fac + nni * f_slv + n_att * f_vec + n_out * f_dky
    return w, dict(F_res=f_res, F_jac=f_jac, F_fac=f_fac, F_slv=f_slv, F_vec=f_vec, F_dky=f_dky, n_res=nre, n_jac=nje,
                   n_fac=nsetups, n_slv=nni, n_att=n_att, n_out=n_out)


# ---- the CPU legs (oracle/ is test + baseline infrastructure: never on the product path) ------------------------------
def cpu_oracle(bundle, fast):
    """fast: the baseline build (static sparse LU, -O3 -march=native, compiled on this machine); else the checker."""
    from oracle.oracle import Oracle
    from pybamm_b200 import build

    o = Oracle(build.build_oracle(bundle, fast=fast))
    o.set_rcm(bundle.jac_rows, bundle.jac_cols)
    o.set_sparse(*build.lu_order(bundle))
    return o


def run_cpu(o, p, n_threads, lu, tol=RTOL):
    t0 = time.perf_counter()
    r = o.solve(T_EVAL, p, t_interp=T_INTERP, rtol=tol, atol=tol, save_vars=True, n_threads=n_threads, lu=lu)
    return r, time.perf_counter() - t0


CPU_NOTE = ("CPU restatement of the reference algorithm (IDA-style BDF, oracle/ida_oracle.c), baseline build: static-pattern "
            "sparse LU in the device's ordering (what klu_refactor does), gcc -O3 -march=native, OpenMP schedule(dynamic) over "
            "systems; not IDAKLU (pybamm / pybammsolvers / casadi / SUNDIALS / KLU are not installable here)")


def _env_ranks(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    return rank, world, local_rank


def run_reference_arm(args):
    """--impl reference: the reference algorithm for this path on the box's host cores, all host threads.  The K timed steps
    together cover the batch exactly once (step k = rows k::K of rank 0's 10 000 input sets) unless that would exceed the
    time budget, in which case a prefix of every slice is taken and `sample` says how many rows ran."""
    rank, world, _ = _env_ranks(args)
    if rank != 0:
        return 0
    from pybamm_b200 import build

    bundle = build.load_bundle(args.model)
    cores = os.cpu_count() or 1
    o = cpu_oracle(bundle, fast=True)
    p_all = sample_inputs(bundle, args.n_sys, np.random.default_rng(20260))
    _, t_cal = run_cpu(o, p_all[:2 * cores], cores, "sparse")          # calibration (also warms the library)
    per_sim = t_cal / (2 * cores)
    steps = max(1, args.steps)
    slices = [np.ascontiguousarray(p_all[k::steps]) for k in range(steps)]
    budget_s = float(args.cpu_budget_s)
    est = per_sim * args.n_sys
    frac = min(1.0, budget_s / max(est, 1e-9))
    if frac < 1.0:
        slices = [s[:max(cores, int(len(s) * frac))] for s in slices]
    for w in range(min(args.warmup, 1)):
        run_cpu(o, slices[0][:4 * cores], cores, "sparse")
    total, n_done = 0.0, 0
    for s in slices:
        _, dt = run_cpu(o, s, cores, "sparse")
        total += dt
        n_done += len(s)
    value = n_done / total
    cfg = workload_config(args, 1, args.n_sys)
    cfg["model_states"] = bundle.n_states
    sample = (f"all {args.n_sys} input sets of rank 0, {steps} timed steps of {len(slices[0])} rows each (rows k::{steps})"
              if frac >= 1.0 else
              f"{n_done} of the {args.n_sys} input sets of rank 0 ({steps} steps x {len(slices[0])} rows; budget {budget_s:.0f} s)")
    line = {
        "impl": "reference", "metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / steps, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "note": CPU_NOTE},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_install": reference_install_probe(),
    }
    print(json.dumps(line), flush=True)
    return 0


def parity_block(bundle, p, n_t, flags, v_gpu, t_end_gpu, n_par, cores):
    """GPU vs the checker oracle at 1e-8 (banded LU, partial pivoting) and vs the oracle at 1e-11 ("truth"), per system."""
    o = cpu_oracle(bundle, fast=False)
    # the pivoting banded LU is the independent checker; with a dense thermal row and column the band is the whole matrix
    # (3523 x 3523 per system), so the largest model is checked against the static sparse LU in the device's ordering
    ref, _ = run_cpu(o, p[:n_par], cores, "band" if bundle.n_states <= 2048 else "sparse")
    tru, _ = run_cpu(o, p[:n_par], cores, "sparse", tol=1e-11)
    # 1e-11 is beyond what the first step's error test accepts for a few input sets (the algebraic states move at O(1) V/s
    # right after the current step; IDA's first-order predictor for them costs h <= rtol |y| / |dy/dt|): those fall back to
    # 1e-10, then 1e-9
    for tol in (1e-10, 1e-9):
        redo = np.nonzero(tru["flag"] < 0)[0]
        if not len(redo):
            break
        again, _ = run_cpu(o, np.ascontiguousarray(p[:n_par][redo]), cores, "sparse", tol=tol)
        for k, i in enumerate(redo):
            for key in ("t", "y"):
                tru[key][i] = again[key][k]
            tru["flag"][i] = again["flag"][k]
    e_gr, e_gt, e_rt, dt_gr, dt_gt = (np.zeros(n_par) for _ in range(5))
    same = True
    for i in range(n_par):
        k = int(n_t[i])
        if k != len(ref["t"][i]) or k != len(tru["t"][i]) or int(flags[i]) != int(ref["flag"][i]) or int(flags[i]) != int(tru["flag"][i]):
            same = False
            e_gr[i] = e_gt[i] = np.inf
            continue
        m = k - 1 if flags[i] == 2 else k          # the event point has its own time in every run
        g, r, t = v_gpu[i, :m], ref["y"][i][:m, 0], tru["y"][i][:m, 0]
        e_gr[i], e_gt[i], e_rt[i] = np.max(np.abs(g - r)), np.max(np.abs(g - t)), np.max(np.abs(r - t))
        dt_gr[i] = abs(t_end_gpu[i] - ref["t"][i][-1])
        dt_gt[i] = abs(t_end_gpu[i] - tru["t"][i][-1])
    # bars: event times within 1e-3 s; voltage within 1e-6 V of the reference algorithm's solution, or,
    # where the 1e-8 tolerances themselves do not deliver 1e-6 V (the last grid point before a cut-off, |dV/dt| ~ 0.1 V/s),
    # no further from the 1e-11 truth than twice the reference algorithm's own error there
    ok_v = (e_gr <= V_BAR) | (e_gt <= np.maximum(V_BAR, 2.0 * e_rt))
    ok = bool(same and ok_v.all() and dt_gr.max() <= T_BAR and dt_gt.max() <= T_BAR)
    return {"systems": n_par, "same_termination": same, "ok": ok,
            "max_abs_dV_gpu_vs_oracle_1e-8": float(e_gr.max()), "max_abs_dV_gpu_vs_truth_1e-11": float(e_gt.max()),
            "max_abs_dV_oracle_1e-8_vs_truth": float(e_rt.max()),
            "systems_above_1e-6_V": {"gpu_vs_oracle": int((e_gr > V_BAR).sum()), "gpu_vs_truth": int((e_gt > V_BAR).sum()),
                                     "oracle_vs_truth": int((e_rt > V_BAR).sum())},
            "median_abs_dV_gpu_vs_truth": float(np.median(e_gt)), "median_abs_dV_oracle_vs_truth": float(np.median(e_rt)),
            "max_abs_dt_end_gpu_vs_oracle": float(dt_gr.max()), "max_abs_dt_end_gpu_vs_truth": float(dt_gt.max()),
            "systems_failing_the_bar": int((~ok_v).sum()),
            "bars": {"dV": V_BAR, "dt": T_BAR,
                     "rule": "per system: |V_gpu - V_oracle(1e-8)| <= 1e-6 V, or |V_gpu - V_truth| <= max(1e-6 V, 2 |V_oracle(1e-8) - V_truth|)"}}, ref


def run_b200_arm(args):
    import torch

    rank, world, local_rank = _env_ranks(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this solver has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        # rank 0 prints ONE line on stdout, the JSON: NCCL's "NCCL version ..." banner (written to file descriptor 1 when the
        # first communicator comes up) goes to stderr instead
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            warm = torch.zeros(1, device=torch.device("cuda", local_rank))
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)

    from pybamm_b200 import BatchedIDASolver, build
    from pybamm_b200.host import Result, canonical_time_grids
    from pybamm_b200.shard import launch_order, partition

    bundle = build.load_bundle(args.model)
    if args.scaling == "strong":
        # the metric's own batch: args.n_sys input sets in TOTAL, dealt to the ranks by shard.partition
        p_all = sample_inputs(bundle, args.n_sys, np.random.default_rng(20260))
        rows = partition(args.n_sys, world)[rank]
        p = np.ascontiguousarray(p_all[rows])
        n_total = args.n_sys
    else:
        p = sample_inputs(bundle, args.n_sys, np.random.default_rng(20260 + rank))
        n_total = args.n_sys * world
    n = len(p)
    solver = BatchedIDASolver(bundle, rtol=RTOL, atol=ATOL, output_variables=["Voltage [V]"], device=local_rank,
                              on_failure="ignore", reuse_buffers=True)
    solver._ensure()
    lib, D = solver._lib, solver.desc
    te, ti = canonical_time_grids(T_EVAL, T_INTERP)
    dev = torch.device("cuda", local_rank)
    d_te = torch.tensor(te, dtype=torch.float64, device=dev)
    d_ti = torch.tensor(ti, dtype=torch.float64, device=dev)
    d_p = torch.tensor(p, dtype=torch.float64, device=dev)
    n_max = lib.pbgpu_result_capacity(len(te), len(ti))
    d_t = torch.zeros((n, n_max), dtype=torch.float64, device=dev)
    d_y = torch.zeros((n, n_max, D.n_vars), dtype=torch.float64, device=dev)
    d_nt = torch.zeros(n, dtype=torch.int32, device=dev)
    d_flag = torch.zeros(n, dtype=torch.int32, device=dev)
    d_stats = torch.zeros((n, 8), dtype=torch.int64, device=dev)
    res = Result()
    res.n_max, res.n_ret = n_max, D.n_vars
    res.t, res.y, res.n_t, res.flag, res.stats = d_t.data_ptr(), d_y.data_ptr(), d_nt.data_ptr(), d_flag.data_ptr(), d_stats.data_ptr()
    stream = torch.cuda.Stream(device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    vp = C.c_void_p
    # launch order of the persistent kernel: expensive rows first (shard.launch_order; results stay in row order).  The cost
    # model is fitted on the step counts of one untimed solve of a 512-row sample (model-free: shard.CostModel)
    solver.solve_arrays(T_EVAL, p[:512], t_interp=T_INTERP)
    order = np.ascontiguousarray(launch_order(p, solver._cost_model), dtype=np.int32)
    if lib.pbgpu_solver_set_launch_order(solver._solver, order.ctypes.data_as(vp), n):
        raise RuntimeError(lib.pbgpu_last_error().decode())

    def launch():
        rc = lib.pbgpu_solve_device(solver._solver, n, vp(d_te.data_ptr()), len(te), vp(d_ti.data_ptr()), len(ti), None, None,
                                    vp(d_p.data_ptr()), 1, C.byref(res), vp(stream.cuda_stream), 0)
        if rc:
            raise RuntimeError(lib.pbgpu_last_error().decode())

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            flush.zero_()
            launch()
        barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        evs = []
        for _ in range(args.steps):
            flush.zero_()                      # L2 flush, outside the event pair
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            launch()
            e1.record(stream)
            evs.append((e0, e1))
        barrier()
        clocks = sampler.stop()
    step_ms = [a.elapsed_time(b) for a, b in evs]
    total_ms = float(sum(step_ms))
    flags = d_flag.cpu().numpy()
    n_t = d_nt.cpu().numpy()
    ok = bool(np.all((flags == 1) | (flags == 2)))
    stats_gpu = d_stats.cpu().numpy().astype(np.float64).mean(axis=0)

    # ---- second number: the same batch returning the FULL state at every output point, device-resident
    # (n x n_max x N doubles: 28 GB for 10 000 DFN systems -- it stays in HBM; one warm-up launch, one timed launch)
    full_state = None
    if world == 1 and not args.no_full_state:
        try:
            d_yf = torch.empty((n, n_max, D.n_states), dtype=torch.float64, device=dev)
            resf = Result()
            resf.n_max, resf.n_ret = n_max, D.n_states
            resf.t, resf.y, resf.n_t, resf.flag, resf.stats = (d_t.data_ptr(), d_yf.data_ptr(), d_nt.data_ptr(),
                                                                d_flag.data_ptr(), d_stats.data_ptr())
            with torch.cuda.stream(stream):
                ms = []
                for k in range(2):
                    flush.zero_()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream)
                    if lib.pbgpu_solve_device(solver._solver, n, vp(d_te.data_ptr()), len(te), vp(d_ti.data_ptr()), len(ti), None, None,
                                              vp(d_p.data_ptr()), 0, C.byref(resf), vp(stream.cuda_stream), 0):
                        raise RuntimeError(lib.pbgpu_last_error().decode())
                    e1.record(stream)
                    torch.cuda.synchronize(dev)
                    ms.append(e0.elapsed_time(e1))
            v_col = d_yf[:64, :, :].cpu().numpy()
            full_state = {"value": n / (ms[-1] * 1e-3), "unit": UNIT, "ms": ms[-1], "hbm_bytes_written": int(8 * (D.n_states + 1) * int(n_t.sum())),
                          "finite": bool(all(np.all(np.isfinite(v_col[i, :int(n_t[i])])) for i in range(len(v_col)))),
                          "note": "output_mode 0: y[n_sys][n_t][n_states] written to HBM, one timed launch after one warm-up"}
            del d_yf, v_col
            torch.cuda.empty_cache()
        except Exception as e:   # e.g. not enough device memory for the full-state block
            full_state = {"value": None, "error": str(e)[:200]}

    # ---- e2e: the public call with host buffers (list of input dicts in, list of Solutions out) -------------------
    inputs_list = [dict(zip(bundle.input_names, (float(x) for x in row))) for row in p]
    sols = solver.solve(T_EVAL, inputs=inputs_list, t_interp=T_INTERP)   # warm-up (pinned buffers, device buffers)
    e2e_steps = max(1, min(args.steps, 3))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        sols = solver.solve(T_EVAL, inputs=inputs_list, t_interp=T_INTERP)
        _ = float(sols[-1]["Voltage [V]"][-1, 0])
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    e2e_kernel_ms = solver.last_kernel_ms
    h2d = 8 * (p.size + len(te) + len(ti))
    d2h = n * (8 * n_max * (1 + D.n_vars) + 4 + 4 + 8 * D.n_states + 64 + 4 + 8)

    # ---- e2e with the full state brought back to the host (what a default solve() user pays), on a bounded sub-batch:
    # 10 000 DFN systems are 28 GB of host memory, so the first --full-state-e2e rows are timed (2.8 GB for 1 000)
    e2e_full = None
    n_fs = min(args.full_state_e2e, n)
    if world == 1 and n_fs > 0 and not args.no_full_state:
        try:
            fs = BatchedIDASolver(bundle, rtol=RTOL, atol=ATOL, device=local_rank, on_failure="ignore", reuse_buffers=True)
            fs._cost_model = solver._cost_model
            fs.solve(T_EVAL, inputs=inputs_list[:n_fs], t_interp=T_INTERP)
            t0 = time.perf_counter()
            sl = fs.solve(T_EVAL, inputs=inputs_list[:n_fs], t_interp=T_INTERP)
            _ = float(sl[-1].y[0, -1])
            dt_fs = time.perf_counter() - t0
            e2e_full = {"value": n_fs / dt_fs, "unit": UNIT, "n_sys": n_fs, "kernel_ms": fs.last_kernel_ms,
                        "d2h_bytes": int(n_fs * 8 * n_max * (1 + D.n_states)),
                        "call": "BatchedIDASolver(output_variables=None).solve(...): y (n_states x n_t) of every system on the host"}
            fs.close()
            del fs, sl
        except Exception as e:
            e2e_full = {"value": None, "error": str(e)[:200]}

    if dist is not None:
        tt = torch.tensor([total_ms, e2e_s, 0.0 if ok else 1.0], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms, e2e_s, bad = float(tt[0]), float(tt[1]), float(tt[2])
        ok = bad == 0.0
    value = n_total * args.steps / (total_ms * 1e-3)
    e2e_value = n_total * e2e_steps / e2e_s

    rc = 0
    if rank == 0:
        cfg = workload_config(args, world, n_total)
        cfg["model_states"] = bundle.n_states
        big = D.smem_bytes > 48 * 1024
        cfg["rows_on_rank_0"] = n
        cfg["waves_per_gpu"] = round(n / (148 * (1 if bundle.n_states > 2048 else 2)), 2) if big else None
        line = {
            "metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": cfg, "clocks": clocks, "gpu_launches": args.steps,   # pbgpu_solve_kernel launches inside the timed region of `value`
            "all_integrations_succeeded": ok,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "kernel_ms_last": e2e_kernel_ms,
                    "call": "BatchedIDASolver.solve(t_eval, inputs=[{...}]*n, t_interp) -> SolutionBatch (list-like; Solution views made on access)"},
            "e2e_full_state": e2e_full,
            "full_state": full_state,
            "gpu_ida_counters_mean": dict(zip(("nsteps", "nrevals", "nlinsetups", "netfails", "nniters", "nncfails", "njevals", "nge"),
                                              stats_gpu.tolist())),
            "reference_install": reference_install_probe(),
        }
        # ---- roofline + CPU baseline + parity (rank 0) ----------------------------------------------------
        cores = os.cpu_count() or 1
        peak_tf = measure_fp64_peak(local_rank)
        n_cpu = min(args.cpu_sample, n) if world == 1 else min(4 * cores, n)
        base, cpu_s = run_cpu(cpu_oracle(bundle, fast=True), p[:n_cpu], cores, "sparse")
        stats_ref = base["stats"].astype(np.float64).mean(axis=0)
        w, terms = work_model(D, stats_ref, float(np.mean(base["n_out"])), bundle.n_states)
        kernel_s = (total_ms / args.steps) * 1e-3
        achieved_tf = w * n / kernel_s / 1e12
        alg_bytes = 8.0 * (p.shape[1] + float(np.mean(n_t)) * (1 + D.n_vars)) + 8
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        traffic, traffic_src = ncu_dram_traffic(n) if args.model == "dfn" else (None, "no capture for this model")
        line["roofline"] = {
            "bound": "fp64", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved_tf / peak_tf,
            "traffic": traffic, "traffic_source": traffic_src, "kernel": "pbgpu_solve_kernel", "flop_per_simulation": w,
            "work_terms": terms,
            "peak_source": "DFMA microbenchmark measured in this run (pbgpu_peak.cu); MEASURED_PEAKS.json has no FP64 entry",
            "hbm": {"achieved": alg_bytes * n / kernel_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": alg_bytes * n / kernel_s / 1e9 / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)",
                    "bytes_per_simulation": alg_bytes},
        }
        n_par = min(args.parity_sample, n) if world == 1 else min(2 * cores, n)
        y_host, t_host = d_y[:n_par].cpu().numpy(), d_t[:n_par].cpu().numpy()
        t_end = np.array([t_host[i, int(n_t[i]) - 1] for i in range(n_par)])
        par, _ = parity_block(bundle, p, n_t, flags, y_host[:, :, 0], t_end, n_par, cores)
        line["parity"] = par
        line["parity_ok"] = bool(par["ok"] and ok)
        if world == 1:
            line["cpu_baseline"] = {"value": n_cpu / cpu_s, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"first {n_cpu} of the {n} input sets, one pass", "note": CPU_NOTE}
        print(json.dumps(line), flush=True)
        if not line["parity_ok"]:
            print("bench.py: PARITY BAR EXCEEDED (see `parity` in the line above)", file=sys.stderr, flush=True)
            rc = 3
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return rc


def run_inproc_arm(args):
    """What north_star describes literally: ONE process, ONE BatchedIDASolver(devices=range(N)).solve(inputs=[...]) call; the
    batch (args.n_sys sets in total) is sharded over the N GPUs by host threads, results gathered on the host."""
    import torch

    from pybamm_b200 import BatchedIDASolver, build

    n_dev = args.inproc_devices
    if torch.cuda.device_count() < n_dev:
        raise SystemExit(f"bench.py: --inproc-devices {n_dev} but only {torch.cuda.device_count()} CUDA devices")
    bundle = build.load_bundle(args.model)
    p = sample_inputs(bundle, args.n_sys, np.random.default_rng(20260))
    inputs_list = [dict(zip(bundle.input_names, (float(x) for x in row))) for row in p]
    solver = BatchedIDASolver(bundle, rtol=RTOL, atol=ATOL, output_variables=["Voltage [V]"], devices=list(range(n_dev)),
                              on_failure="ignore", reuse_buffers=True)
    for _ in range(max(1, min(args.warmup, 2))):
        sols = solver.solve(T_EVAL, inputs=inputs_list, t_interp=T_INTERP)
    steps = max(1, min(args.steps, 5))
    sampler = ClockSampler(0)
    sampler.start()
    t0 = time.perf_counter()
    for _ in range(steps):
        sols = solver.solve(T_EVAL, inputs=inputs_list, t_interp=T_INTERP)
        _ = float(sols[-1]["Voltage [V]"][-1, 0])
    dt = time.perf_counter() - t0
    clocks = sampler.stop()
    cfg = workload_config(args, n_dev, args.n_sys)
    cfg["model_states"] = bundle.n_states
    cfg["sharding"] = f"one process, {n_dev} device(s), one host thread per device (BatchedIDASolver(devices=...))"
    value = args.n_sys * steps / dt
    line = {"metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": n_dev, "steps": steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": cfg, "clocks": clocks, "gpu_launches": steps * n_dev, "impl": "b200-inproc",
            "e2e": {"value": value, "unit": UNIT, "kernel_ms_last_max_over_devices": solver.last_kernel_ms,
                    "call": f"BatchedIDASolver(devices=range({n_dev})).solve(t_eval, inputs=[{{...}}]*{args.n_sys}, t_interp)"},
            "all_integrations_succeeded": bool(all(s.flag in (1, 2) for s in sols))}
    print(json.dumps(line), flush=True)
    return 0


def sample_inputs_c5(bundle, n, rng):
    """Config 5: the inputs of config 3 minus the current; the schedule's own inputs are set by the schedule."""
    p = np.tile(np.asarray(bundle.meta["nominal_inputs"], dtype=np.float64), (n, 1))
    for nm, fn in (("Negative particle diffusivity [m2.s-1]", lambda: 10 ** rng.uniform(-14, -13, n)),
                   ("Positive particle diffusivity [m2.s-1]", lambda: 10 ** rng.uniform(-15, -14, n)),
                   ("Initial concentration in electrolyte [mol.m-3]", lambda: rng.uniform(900, 1100, n))):
        p[:, bundle.input_names.index(nm)] = fn()
    return np.ascontiguousarray(p)


def run_experiment_arm(args):
    """BASELINE config 5: DFN unified-experiment model, 3 x (discharge 1C to 2.5 V / rest 1 h / charge 0.3C to 4.2 V / hold 4.2 V
    to C/50), n_sys input sets, the whole schedule inside one kernel launch per step of the bench.  The reference has no
    batched path for this (Simulation.solve refuses a list of inputs with an experiment, simulation.py:832-835): its baseline is
    a loop over the input sets of the per-step solver.step loop -- here that loop around the CPU oracle, one thread per set."""
    import torch
    from concurrent.futures import ThreadPoolExecutor

    from oracle.experiment_loop import oracle_experiment
    from pybamm_b200 import BatchedIDASolver, build
    from pybamm_b200.experiment import ExperimentSchedule

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this solver has no CPU fallback")
    bundle = build.load_bundle(args.model)
    sched = ExperimentSchedule.from_meta(bundle.meta["experiment"])
    n = args.n_sys
    p = sample_inputs_c5(bundle, n, np.random.default_rng(20260))
    solver = BatchedIDASolver(bundle, rtol=RTOL, atol=ATOL, output_variables=["Voltage [V]", "Current [A]"], on_failure="ignore")
    for _ in range(max(1, min(args.warmup, 2))):
        r = solver.solve_experiment(sched, p)
    steps = max(1, min(args.steps, 5))
    sampler = ClockSampler(0)
    sampler.start()
    t0 = time.perf_counter()
    kernel_ms = []
    for _ in range(steps):
        r = solver.solve_experiment(sched, p)
        kernel_ms.append(solver.last_kernel_ms)
    dt = time.perf_counter() - t0
    clocks = sampler.stop()
    ok = bool(np.all(r["flag"] == 1))
    cores = os.cpu_count() or 1
    o = cpu_oracle(bundle, fast=True)
    n_cpu = min(2 * cores, n)

    def one(i):
        return oracle_experiment(bundle, o, sched, p[i:i + 1], rtol=RTOL, atol=ATOL, save_vars=False, lu="sparse")[0]

    tc = time.perf_counter()
    with ThreadPoolExecutor(max_workers=cores) as pool:
        refs = list(pool.map(one, range(n_cpu)))
    cpu_s = time.perf_counter() - tc
    # parity on the CPU sample: same step terminations and point counts, end voltages of every step, voltage over the grid
    worst_v, worst_t, same = 0.0, 0.0, True
    for i, ref in enumerate(refs):
        if not (np.array_equal(ref["step_end"], r["step_end"][i]) and np.array_equal(ref["step_term"], r["step_term"][i])):
            same = False
            continue
        k = int(r["n_t"][i])
        ends = set(ref["step_end"] - 1)
        inner = np.array([j for j in range(k) if j not in ends])
        v_ref = np.array([bundle.eval_variable("Voltage [V]", ref["t"][j], ref["y"][j], np.r_[p[i]])[0] for j in inner[::7]])
        worst_v = max(worst_v, float(np.max(np.abs(r["y"][i, inner[::7], 0] - v_ref))))
        worst_t = max(worst_t, float(np.max(np.abs(r["t"][i, :k] - ref["t"]))))
    cfg = {"workload": f"DFN Chen2020 unified experiment model, CCCV x 3 cycles (12 steps), {n} input sets varying 3 parameters",
           "model_states": bundle.n_states, "n_sys_total": n, "rtol": RTOL, "atol": ATOL, "steps_of_the_schedule": sched.n_steps,
           "output": "Voltage [V], Current [A] at every step's period grid", "l2": "inputs larger than L2 are not needed: state is on chip"}
    value = n * steps / dt
    line = {"metric": f"DFN CCCV experiments/sec ({n} input sets, 3 cycles)", "value": n / (float(np.mean(kernel_ms)) * 1e-3), "unit": "experiments/s",
            "n_gpus": 1, "steps": steps, "warmup": args.warmup, "ms_per_step": float(np.mean(kernel_ms)), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg, "clocks": clocks,
            "gpu_launches": steps, "all_integrations_succeeded": ok,
            "e2e": {"value": value, "unit": "experiments/s", "call": "BatchedIDASolver.solve_experiment(schedule, inputs) with host buffers",
                    "d2h_bytes_per_step": int(n * sched.capacity * 8 * 3)},
            "gpu_ida_counters_mean": dict(zip(("nsteps", "nrevals", "nlinsetups", "netfails", "nniters", "nncfails", "njevals", "nge"),
                                              r["stats"].astype(float).mean(axis=0).tolist())),
            "cpu_baseline": {"value": n_cpu / cpu_s, "unit": "experiments/s", "cores": cores, "kind": "port",
                             "sample": f"first {n_cpu} input sets, one thread each",
                             "note": "the reference's per-step solver.step loop restated around the CPU oracle (oracle/experiment_loop.py), "
                                     "static sparse LU, -O3 -march=native; the reference itself has no batched experiment path"},
            "parity": {"systems": n_cpu, "same_step_terminations_and_counts": same, "max_abs_dV": worst_v, "max_abs_dt": worst_t,
                       "bars": {"dV": 5e-6, "dt": 5e-2, "why": "hold steps end where |dI/dt| ~ 1e-6 A/s: event times are defined to ~1e-2 s at 1e-8"}},
            }
    line["parity_ok"] = bool(same and ok and worst_v <= 5e-6 and worst_t <= 5e-2)
    print(json.dumps(line), flush=True)
    return 0 if line["parity_ok"] else 3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="dfn", choices=["spm", "spme", "dfn", "dfn_cccv", "dfn_thermal40"])
    ap.add_argument("--n-sys", type=int, default=10000, help="input sets per GPU (weak) or in total (strong, in-process)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--inproc-devices", type=int, default=0, help="run the one-process multi-device arm on this many GPUs")
    ap.add_argument("--cpu-sample", type=int, default=1024, help="systems of the cpu_baseline / work-model sample")
    ap.add_argument("--parity-sample", type=int, default=256, help="systems compared with the oracle at 1e-8 and 1e-11")
    ap.add_argument("--cpu-budget-s", type=float, default=150.0, help="--impl reference: time budget of the timed steps")
    ap.add_argument("--full-state-e2e", type=int, default=1000, help="rows of the full-state end-to-end number (0: skip)")
    ap.add_argument("--no-full-state", action="store_true", help="skip the full-state (output_mode 0) secondary numbers")
    args = ap.parse_args()
    if args.model == "dfn_cccv":
        if args.n_sys == 10000:
            args.n_sys = 4000          # BASELINE config 5
        return run_experiment_arm(args)
    if args.model == "dfn_thermal40":
        args.no_full_state = True      # the full state of 10 000 x 362 points x 3523 states is 102 GB
    if args.impl == "reference":
        return run_reference_arm(args)
    if args.inproc_devices > 0:
        return run_inproc_arm(args)
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
