#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric: DFN simulations / second (10k input sets per GPU, rtol = atol = 1e-8).

  python bench.py --gpus N --steps K --warmup W                (N > 1: launched by torch.distributed.run, one rank per GPU)
  python bench.py --impl reference --gpus N --steps K --warmup W     (CPU arm: the reference algorithm on the host cores)

A "step" is one batched solve of the workload: DFN (Chen2020, default mesh, 962 states), constant-current
discharge, t_eval = [0, 3700] s, t_interp = linspace(0, 3600, 361), 10 000 input sets per GPU varying four
parameters (current, both particle diffusivities, initial electrolyte concentration),
output_variables = ["Voltage [V]"].  The batch is sharded by rank with no collective inside the integration
(weak scaling: every rank integrates its own 10k sets).

`value`  : device-resident inputs, pbgpu_solve_device on a CUDA stream, CUDA events around each step.
`e2e`    : the public call BatchedIDASolver.solve(t_eval, inputs=[{...}, ...]) with host buffers: flattening the
           input dicts, H2D, kernel, D2H and building the Solution objects are all inside the timed region.
`roofline`: FP64 pipe. achieved = W(algorithmic flop per simulation) * sims / kernel time; peak = DFMA
           rate measured live on this box by pybamm_b200/csrc/pbgpu_peak.cu.  The HBM fraction is reported beside it.
`cpu_baseline`: the CPU oracle (oracle/ida_oracle.c, "port": the reference IDAKLU cannot be built here) on all host
           cores over a bounded sample of the same input sets; its IDA counters also feed W.
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
METRIC = "DFN simulations/sec (10k input sets)"   # --model spm/spme rename it below (parity-config side runs)
UNIT = "simulations/s"


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


def workload_config(args, world):
    return {
        "workload": f"{args.model.upper()} Chen2020 default mesh, 1C-range constant-current discharge, "
                    f"{args.n_sys} input sets per GPU varying {4 if args.model == 'dfn' else 2} parameters",
        "model_states": None,  # filled by the caller
        "n_sys_per_gpu": args.n_sys,
        "n_sys_total": args.n_sys * world,
        "rtol": RTOL, "atol": ATOL,
        "t_eval": T_EVAL, "t_interp": "linspace(0, 3600, 361)",
        "output": "Voltage [V]",
        "sharding": f"batch rows over {world} rank(s), no collective",
        "l2": "flushed between timed steps (256 MiB write)",
        "launch_order": "rows handed to the teams by descending predicted step count (log-linear fit on a previous solve), results in row order",
    }


def ncu_dram_traffic(n_sys):
    """dram__bytes_read.sum + dram__bytes_write.sum of pbgpu_solve_kernel from the committed `ncu --set full` capture
    (profiles/r01_dfn_full_ncu_raw.csv: `ncu --page raw --csv` of one launch with one DFN system per team, so the
    grid size is the number of systems), scaled to this launch's system count.
    Returns (bytes per launch or None, provenance)."""
    path = os.path.join(ROOT, "profiles", "r01_dfn_full_ncu_raw.csv")
    try:
        import csv

        rows = [r for r in csv.reader(open(path)) if r]
        hdr, units, vals = rows[0], rows[1], rows[2]
        col = {h: k for k, h in enumerate(hdr)}
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        total = sum(float(vals[col[m]].replace(",", "")) * scale[units[col[m]]]
                    for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        n_cap = int(float(vals[col["launch__grid_size"]].replace(",", "")))
        per_sim = total / n_cap
        return per_sim * n_sys, (f"profiles/r01_dfn_full_ncu_raw.csv: ncu --set full, one launch of {n_cap} systems "
                                 f"({per_sim / 1e3:.1f} KB per simulation), scaled to {n_sys} systems")
    except Exception:
        pass
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


def run_cpu_oracle(bundle, p, n_threads):
    """Reference algorithm on the host cores (the CPU oracle; test/bench infrastructure only)."""
    from oracle.oracle import Oracle
    from pybamm_b200 import build

    o = Oracle(build.build_oracle(bundle))
    o.set_rcm(bundle.jac_rows, bundle.jac_cols)
    t0 = time.perf_counter()
    r = o.solve(T_EVAL, p, t_interp=T_INTERP, rtol=RTOL, atol=ATOL, save_vars=True, n_threads=n_threads)
    return r, time.perf_counter() - t0


def _env_ranks(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    return rank, world, local_rank


def run_reference_arm(args):
    """--impl reference: the reference algorithm for this path on the box's host cores.  pybamm / pybammsolvers / casadi /
    SUNDIALS / KLU cannot be installed or built here, so this runs the CPU oracle ("port") with all host threads
    on a bounded sample of rank 0's input sets per step."""
    rank, world, _ = _env_ranks(args)
    if rank != 0:
        return 0
    from pybamm_b200 import build

    bundle = build.load_bundle(args.model)
    cores = os.cpu_count() or 1
    p_all = sample_inputs(bundle, args.n_sys, np.random.default_rng(20260))
    _, t_cal = run_cpu_oracle(bundle, p_all[:2 * cores], cores)          # calibration (also warms the library)
    per_sim = t_cal / (2 * cores)
    budget_s = 150.0
    n_sample = int(max(cores, min(args.n_sys, budget_s / max(per_sim, 1e-6) / (args.steps + args.warmup))))
    n_sample = min(n_sample, args.cpu_sample_max)
    p = np.ascontiguousarray(p_all[:n_sample])
    for _ in range(args.warmup):
        run_cpu_oracle(bundle, p, cores)
    total = 0.0
    for _ in range(args.steps):
        _, dt = run_cpu_oracle(bundle, p, cores)
        total += dt
    value = n_sample * args.steps / total
    cfg = workload_config(args, 1)
    cfg["model_states"] = bundle.n_states
    sample = f"first {n_sample} of the {args.n_sys} input sets of rank 0, per step"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "note": "CPU oracle (IDA restatement + banded LU), OpenMP over systems; not IDAKLU"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


def run_b200_arm(args):
    import torch

    rank, world, local_rank = _env_ranks(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this solver has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("NCCL_DEBUG", "WARN")   # no "NCCL version" banner on stdout: rank 0 prints ONE line, the JSON
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from pybamm_b200 import BatchedIDASolver, build
    from pybamm_b200.host import Result, canonical_time_grids

    bundle = build.load_bundle(args.model)
    n = args.n_sys
    p = sample_inputs(bundle, n, np.random.default_rng(20260 + rank))
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
    # launch order of the persistent kernel: expensive rows first (shard.launch_order; results stay in row order)
    from pybamm_b200.shard import launch_order

    # the cost model is fitted on the step counts of one untimed solve of a 512-row sample (model-free: shard.CostModel)
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
    d2h = n * (8 * n_max * (1 + D.n_vars) + 4 + 4 + 8 * D.n_states + 64)

    if dist is not None:
        tt = torch.tensor([total_ms, e2e_s, 0.0 if ok else 1.0], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms, e2e_s, bad = float(tt[0]), float(tt[1]), float(tt[2])
        ok = bad == 0.0
    value = n * world * args.steps / (total_ms * 1e-3)
    e2e_value = n * world * e2e_steps / e2e_s

    line = None
    if rank == 0:
        cfg = workload_config(args, world)
        cfg["model_states"] = bundle.n_states
        cfg["teams_per_launch"] = None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": cfg, "clocks": clocks, "gpu_launches": args.steps,   # pbgpu_solve_kernel launches inside the timed region of `value`
            "all_integrations_succeeded": ok,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "kernel_ms_last": e2e_kernel_ms,
                    "call": "BatchedIDASolver.solve(t_eval, inputs=[{...}]*n, t_interp) -> list[Solution]"},
            "full_state": full_state,
            "gpu_ida_counters_mean": dict(zip(("nsteps", "nrevals", "nlinsetups", "netfails", "nniters", "nncfails", "njevals", "nge"),
                                              stats_gpu.tolist())),
        }
        # ---- roofline + CPU baseline (rank 0) -------------------------------------------------------------
        cores = os.cpu_count() or 1
        peak_tf = measure_fp64_peak(local_rank)
        n_cpu = min(args.cpu_sample, n) if world == 1 else min(4 * cores, n)
        ref, cpu_s = run_cpu_oracle(bundle, p[:n_cpu], cores)
        stats_ref = ref["stats"].astype(np.float64).mean(axis=0)
        w, terms = work_model(D, stats_ref, float(np.mean(ref["n_out"])), bundle.n_states)
        kernel_s = (total_ms / args.steps) * 1e-3
        achieved_tf = w * n / kernel_s / 1e12
        alg_bytes = 8.0 * (p.shape[1] + float(np.mean(n_t)) * (1 + D.n_vars)) + 8
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        traffic, traffic_src = ncu_dram_traffic(n)
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
        # parity of the GPU results against the oracle on the sampled systems
        y_host, t_host = d_y[:n_cpu].cpu().numpy(), d_t[:n_cpu].cpu().numpy()
        wv = wt = wk = 0.0
        agree = True
        for i in range(n_cpu):
            k = int(n_t[i])
            if k != len(ref["t"][i]) or int(flags[i]) != int(ref["flag"][i]):
                agree = False
                continue
            m = k - 1 if flags[i] == 2 else k
            dv = np.abs(y_host[i, :m, 0] - ref["y"][i][:m, 0])
            wv = max(wv, float(np.max(dv)))
            # the same without the last grid point before a voltage cut-off: on that knee |dV/dt| is ~100x larger and two
            # integrations at rtol = atol = 1e-8 differ by more than 1e-6 V from EACH OTHER'S tighter-tolerance limit
            wk = max(wk, float(np.max(dv[:m - 1]))) if flags[i] == 2 and m > 1 else max(wk, float(np.max(dv)))
            wt = max(wt, abs(float(t_host[i, k - 1]) - float(ref["t"][i][-1])))
        line["parity_vs_oracle"] = {"systems": n_cpu, "max_abs_dV": wv, "max_abs_dV_before_the_cutoff_knee": wk,
                                    "max_abs_dt_end": wt, "same_termination": agree, "bars": {"dV": 1e-6, "dt": 1e-3}}
        if world == 1:
            line["cpu_baseline"] = {"value": n_cpu / cpu_s, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"first {n_cpu} of the {n} input sets, one pass",
                                    "note": "CPU oracle (IDA restatement + banded LU), OpenMP over systems; not IDAKLU"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="dfn", choices=["spm", "spme", "dfn"])
    ap.add_argument("--n-sys", type=int, default=10000, help="input sets per GPU")
    ap.add_argument("--cpu-sample", type=int, default=1024, help="systems of the cpu_baseline / work-model sample")
    ap.add_argument("--cpu-sample-max", type=int, default=2048)
    ap.add_argument("--no-full-state", action="store_true", help="skip the full-state (output_mode 0) secondary number")
    args = ap.parse_args()
    global METRIC
    if args.model != "dfn":
        METRIC = f"{args.model.upper()} simulations/sec ({args.n_sys} input sets)"
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
