"""Build + drive the host emulation of the device integrator (tests only)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from pybamm_b200 import build
from pybamm_b200.host import Options, LaunchArgs, ObserveArgs, Schedule, default_options, canonical_time_grids

HERE = os.path.dirname(os.path.abspath(__file__))


def build_emu(bundle, force=False):
    header, _ = build.generate(bundle, "cuda", build.BUILD)
    outdir = os.path.join(HERE, "_build")
    os.makedirs(outdir, exist_ok=True)
    out = os.path.join(outdir, f"emu_{bundle.name}.so")
    srcs = [os.path.join(HERE, "emu_main.cpp"), header, os.path.join(build.CSRC, "pbgpu_ida.cuh"),
            os.path.join(build.CSRC, "pbgpu_team.cuh"), os.path.join(build.ROOT, "include", "pbgpu.h")]
    dig = build._src_digest(srcs)
    stamp = out + ".stamp"
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == dig:
        return out
    cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", "-ffp-contract=off",
           f"-DPBM_HEADER=\"{header}\"", f"-DPB_GLOBAL_VEC={1 if bundle.n_states > 256 else 0}", f"-I{build.CSRC}", f"-I{build.ROOT}/include", srcs[0], "-o", out]
    subprocess.run(cmd, check=True)
    with open(stamp, "w") as f:
        f.write(dig)
    return out


class Emu:
    def __init__(self, so):
        self.lib = C.CDLL(so)
        sizes = np.zeros(8, np.int32)
        self.lib.emu_sizes(sizes.ctypes.data_as(C.c_void_p))
        self.n, self.n_diff, self.n_p, self.n_ev, self.nnz, self.n_var, self.nnz_lu, self.smem = (int(x) for x in sizes)

    def eval(self, which, y, p, t=0.0, cj=0.0, y2=None):
        idx = {"res": 0, "jac": 1, "events": 2, "ic": 3, "vars": 4, "solve": 5}[which]
        n_out = {"res": self.n, "jac": self.nnz, "events": max(self.n_ev, 1), "ic": self.n,
                 "vars": max(self.n_var, 1), "solve": self.n}[which]
        out = np.zeros(n_out)
        y = np.ascontiguousarray(y if y is not None else np.zeros(self.n), dtype=np.float64)
        p = np.ascontiguousarray(p, dtype=np.float64)
        y2 = np.ascontiguousarray(y2 if y2 is not None else np.zeros(self.n), dtype=np.float64)
        vp = C.c_void_p
        self.lib.emu_eval.argtypes = [C.c_int, vp, vp, C.c_double, C.c_double, vp, vp]
        rc = self.lib.emu_eval(idx, y.ctypes.data_as(vp), p.ctypes.data_as(vp), float(t), float(cj),
                               y2.ctypes.data_as(vp), out.ctypes.data_as(vp))
        return out, rc

    def solve(self, t_eval, inputs, t_interp=None, rtol=1e-4, atol=1e-6, y0=None, save_vars=False, options=None, n_max=None,
              want_yp=False, yp0=None):
        t_eval, t_interp = canonical_time_grids(t_eval, t_interp)
        inputs = np.ascontiguousarray(np.atleast_2d(np.asarray(inputs, dtype=np.float64)))
        n_sys = inputs.shape[0]
        atol_v = np.ascontiguousarray(np.broadcast_to(np.asarray(atol, dtype=np.float64), (self.n,)).copy())
        n_ret = self.n_var if save_vars else self.n
        if n_max is None:
            n_max = (len(t_interp) + len(t_eval) + 2) if len(t_interp) else 20000
        t_out = np.zeros((n_sys, n_max)); y_out = np.zeros((n_sys, n_max, n_ret))
        n_out = np.zeros(n_sys, np.int32); flag = np.zeros(n_sys, np.int32)
        y_term = np.zeros((n_sys, self.n)); stats = np.zeros((n_sys, 8), np.int64)
        event_idx = np.full(n_sys, -7, np.int32); t_event = np.zeros(n_sys)
        yp_out = np.zeros((n_sys, n_max, self.n)) if (want_yp and not save_vars) else None
        queue = np.zeros(1, np.uint32)
        opt = default_options(rtol=rtol, **(options or {}))
        L = LaunchArgs()
        L.n_sys, L.n_eval, L.n_interp, L.n_max, L.save_vars, L.n_ret = n_sys, len(t_eval), len(t_interp), n_max, int(save_vars), n_ret
        vp = C.c_void_p
        keep = [t_eval, t_interp, inputs, atol_v]
        L.t_eval = t_eval.ctypes.data_as(vp); L.t_interp = t_interp.ctypes.data_as(vp) if len(t_interp) else None
        L.inputs = inputs.ctypes.data_as(vp)
        if y0 is not None:
            y0 = np.ascontiguousarray(np.atleast_2d(np.asarray(y0, dtype=np.float64))); keep.append(y0)
            L.y0 = y0.ctypes.data_as(vp)
        else:
            L.y0 = None
        L.atol = atol_v.ctypes.data_as(vp)
        L.t_out = t_out.ctypes.data_as(vp); L.y_out = y_out.ctypes.data_as(vp); L.n_out = n_out.ctypes.data_as(vp)
        L.flag = flag.ctypes.data_as(vp); L.y_term = y_term.ctypes.data_as(vp); L.stats = stats.ctypes.data_as(vp)
        L.queue = queue.ctypes.data_as(vp); L.opt = opt
        L.event_idx = event_idx.ctypes.data_as(vp); L.t_event = t_event.ctypes.data_as(vp)
        L.yp_out = yp_out.ctypes.data_as(vp) if yp_out is not None else None
        if yp0 is not None:
            yp0 = np.ascontiguousarray(np.atleast_2d(np.asarray(yp0, dtype=np.float64))); keep.append(yp0)
            L.yp0 = yp0.ctypes.data_as(vp)
        self.lib.emu_solve(C.byref(L))
        ns = np.minimum(n_out, n_max)
        return dict(t=[t_out[i, :ns[i]] for i in range(n_sys)], y=[y_out[i, :ns[i]] for i in range(n_sys)],
                    yp=None if yp_out is None else [yp_out[i, :ns[i]] for i in range(n_sys)],
                    flag=flag, y_term=y_term, stats=stats, n_out=n_out, event_idx=event_idx, t_event=t_event)

    def observe(self, t, y, n_t, inputs, yp=None, t_obs=None):
        """The device observation routine (pb_observe_system) on the host. Padded arrays as BatchedIDASolver.observe."""
        t = np.ascontiguousarray(t, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        n_sys, n_max = t.shape
        n_t = np.ascontiguousarray(n_t, dtype=np.int32)
        p = np.ascontiguousarray(np.asarray(inputs, dtype=np.float64).reshape(n_sys, -1))
        if p.shape[1] == 0:
            p = np.zeros((n_sys, 1))
        rows = n_max if t_obs is None else len(t_obs)
        out = np.zeros((n_sys, rows, self.n_var))
        O = ObserveArgs()
        vp = C.c_void_p
        O.n_sys, O.n_max, O.n_obs = n_sys, n_max, 0 if t_obs is None else len(t_obs)
        O.n_t, O.t, O.y, O.inputs, O.out = (a.ctypes.data_as(vp) for a in (n_t, t, y, p, out))
        keep = []
        if yp is not None:
            yp = np.ascontiguousarray(yp, dtype=np.float64); keep.append(yp); O.yp = yp.ctypes.data_as(vp)
        if t_obs is not None:
            t_obs = np.ascontiguousarray(t_obs, dtype=np.float64); keep.append(t_obs); O.t_obs = t_obs.ctypes.data_as(vp)
        if self.lib.emu_observe(C.byref(O)):
            raise RuntimeError("the model has no compiled output variables")
        return out

    def solve_experiment(self, bundle, schedule, inputs, t0=0.0, rtol=1e-4, atol=1e-6, y0=None, save_vars=False, options=None):
        """pb_solve_experiment (the device experiment loop) on the host."""
        inputs = np.ascontiguousarray(np.atleast_2d(np.asarray(inputs, dtype=np.float64)))
        n_sys = inputs.shape[0]
        atol_v = np.ascontiguousarray(np.broadcast_to(np.asarray(atol, dtype=np.float64), (self.n,)).copy())
        n_ret = self.n_var if save_vars else self.n
        n_max = schedule.capacity
        ns = schedule.n_steps
        t_out = np.zeros((n_sys, n_max)); y_out = np.zeros((n_sys, n_max, n_ret))
        n_out = np.zeros(n_sys, np.int32); flag = np.zeros(n_sys, np.int32)
        y_term = np.zeros((n_sys, self.n)); stats = np.zeros((n_sys, 8), np.int64)
        event_idx = np.full(n_sys, -7, np.int32); t_event = np.zeros(n_sys)
        step_end = np.zeros((n_sys, ns), np.int32); step_term = np.zeros((n_sys, ns), np.int32)
        queue = np.zeros(1, np.uint32)
        slots = schedule.slots(bundle)
        vp = C.c_void_p
        L = LaunchArgs()
        L.n_sys, L.n_eval, L.n_interp, L.n_max, L.save_vars, L.n_ret = n_sys, 2, 0, n_max, int(save_vars), n_ret
        L.inputs = inputs.ctypes.data_as(vp)
        if y0 is not None:
            y0 = np.ascontiguousarray(np.atleast_2d(np.asarray(y0, dtype=np.float64)))
            L.y0 = y0.ctypes.data_as(vp)
        L.atol = atol_v.ctypes.data_as(vp)
        L.t_out = t_out.ctypes.data_as(vp); L.y_out = y_out.ctypes.data_as(vp); L.n_out = n_out.ctypes.data_as(vp)
        L.flag = flag.ctypes.data_as(vp); L.y_term = y_term.ctypes.data_as(vp); L.stats = stats.ctypes.data_as(vp)
        L.event_idx = event_idx.ctypes.data_as(vp); L.t_event = t_event.ctypes.data_as(vp)
        L.queue = queue.ctypes.data_as(vp); L.opt = default_options(rtol=rtol, **(options or {}))
        q = Schedule()
        q.n_steps, q.n_over, q.experiment_event = ns, len(slots), schedule.event_index(bundle)
        q.over_slot = slots.ctypes.data_as(vp); q.over_val = schedule.over_val.ctypes.data_as(vp)
        q.duration = schedule.duration.ctypes.data_as(vp); q.n_pts = schedule.n_pts.ctypes.data_as(vp)
        q.stop_on_final_time = schedule.stop_on_final_time.ctypes.data_as(vp)
        L.sched = q; L.t_start = float(t0)
        L.step_end = step_end.ctypes.data_as(vp); L.step_term = step_term.ctypes.data_as(vp)
        self.lib.emu_solve(C.byref(L))
        return dict(t=[t_out[i, :n_out[i]] for i in range(n_sys)], y=[y_out[i, :n_out[i]] for i in range(n_sys)], flag=flag,
                    y_term=y_term, stats=stats, n_out=n_out, event_idx=event_idx, t_event=t_event, step_end=step_end,
                    step_term=step_term)
