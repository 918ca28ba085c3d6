"""Host logic and the drop-in boundary on the CPU tier (no compute calls without a GPU)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from pybamm_b200 import BatchedIDASolver, SolverError, build, shard
from pybamm_b200.host import canonical_time_grids, default_options, OPTION_DEFAULTS
from tests import util

ROOT = util.ROOT


def _declared_functions():
    text = open(os.path.join(ROOT, "include", "pbgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pbgpu_[a-z_]+)\s*\(", text)))


def test_c_abi_library_exports_every_declared_symbol():
    names = _declared_functions()
    assert {"pbgpu_model_load", "pbgpu_solver_create", "pbgpu_solve", "pbgpu_solve_device", "pbgpu_model_eval", "pbgpu_solver_set_launch_order",
            "pbgpu_last_error", "pbgpu_result_capacity", "pbgpu_default_options"} <= set(names)
    lib = ctypes.CDLL(build.build_api())
    for n in names:
        assert hasattr(lib, n), f"libpbgpu.so does not export {n}"


def test_model_images_export_the_image_interface():
    for name in util.BATTERY:
        path = os.path.join(build.BUILD, f"pbgpu_model_{name}.so")
        if not os.path.exists(path):
            path = build.build_model(util.bundle(name))
        syms = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True).stdout
        for s in ("pbgpu_image_describe", "pbgpu_image_launch", "pbgpu_image_eval"):
            assert s in syms
        sass = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", path], capture_output=True, text=True).stdout
        assert "sm_100a" in sass


def test_default_options_match_reference_defaults():
    # packages/pybamm/src/pybamm/solvers/idaklu_solver.py:286-325
    lib = ctypes.CDLL(build.build_api())
    from pybamm_b200.host import Options

    o = Options()
    lib.pbgpu_default_options(ctypes.byref(o))
    py = default_options()
    for k, v in OPTION_DEFAULTS.items():
        assert getattr(o, k) == type(getattr(o, k))(v), k
        assert getattr(py, k) == getattr(o, k), k
    assert o.max_order_bdf == 5 and o.max_num_steps == 100000 and o.max_nonlinear_iterations == 40
    assert o.max_convergence_failures == 100 and o.nonlinear_convergence_coefficient == 0.33


def test_result_capacity():
    lib = ctypes.CDLL(build.build_api())
    assert lib.pbgpu_result_capacity(2, 361) >= 2 + 361


def test_time_grid_canonicalisation():
    # idaklu_source/IDAKLUSolverGroup.cpp:25-43
    te, ti = canonical_time_grids([3.0, 0.0, 3.0, 1.0], [0.5, 1.0, 0.5, 2.0, 0.0])
    np.testing.assert_array_equal(te, [0.0, 1.0, 3.0])
    np.testing.assert_array_equal(ti, [0.5, 2.0])
    with pytest.raises(ValueError, match="at least 2"):
        canonical_time_grids([1.0], None)
    with pytest.raises(ValueError, match="greater than the smallest"):
        canonical_time_grids([0.0, 1.0], [-0.5])
    with pytest.raises(ValueError, match="less than the greatest"):
        canonical_time_grids([0.0, 1.0], [1.5])


def test_input_flattening_follows_sorted_names():
    b = util.bundle("kat_inputs")   # a (scalar), b (size 2): base_solver.py:1824, idaklu_solver.py:19-23
    p = b.flatten_inputs([{"b": np.array([[0.2], [0.3]]), "a": 0.1}, {"a": 1.0, "b": [2.0, 3.0]}])
    np.testing.assert_array_equal(p, [[0.1, 0.2, 0.3], [1.0, 2.0, 3.0]])
    with pytest.raises(ValueError, match="No value provided for input"):
        b.flatten_inputs([{"a": 0.1}])


def test_solver_argument_errors_and_loud_failure_without_gpu():
    b = util.bundle("spm")
    with pytest.raises(SolverError, match="unknown IDAKLU option"):
        BatchedIDASolver(b, options={"no_such_option": 1})
    with pytest.raises(SolverError, match="less than or equal to the number of threads"):
        BatchedIDASolver(b, options={"num_threads": 2, "num_solvers": 4})   # idaklu_source/Options.cpp:29-33
    with pytest.raises(SolverError, match="not compiled into model bundle"):
        BatchedIDASolver(b, output_variables=["Nope"])
    s = BatchedIDASolver(b, options={"num_threads": 8, "print_stats": False})   # accepted and ignored
    import torch

    if not torch.cuda.is_available():
        # no CPU fallback: the product path fails loudly when there is no usable CUDA device
        with pytest.raises(SolverError, match="no usable CUDA device|CPU fallback|cannot load"):
            s.solve_arrays([0, 1], np.zeros((1, b.n_inputs)))


def test_product_package_never_imports_the_oracle():
    for root, _, files in os.walk(os.path.join(ROOT, "pybamm_b200")):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h")):
                text = open(os.path.join(root, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "tests.emu" not in text, f


def test_partition_covers_and_balances():
    for n, w in ((10, 3), (7, 8), (1000, 8), (5, 1)):
        rng = np.random.default_rng(n)
        cost = rng.uniform(1, 5, n)
        parts = shard.partition(n, w, cost)
        allidx = np.concatenate(parts)
        assert sorted(allidx.tolist()) == list(range(n))
        sizes = [len(p) for p in parts]
        assert max(sizes) - min(sizes) <= 1
        assert all(np.all(np.diff(p) > 0) for p in parts if len(p) > 1)
        if n >= 100:
            sums = [cost[p].sum() for p in parts]
            assert (max(sums) - min(sums)) / np.mean(sums) < 0.05
    parts = shard.partition(9, 2)
    vals = [np.arange(9.0)[p][:, None] for p in parts]
    np.testing.assert_array_equal(shard.gather_rows(vals, parts, 9)[:, 0], np.arange(9.0))


_GLOO_WORKER = r'''
import os, sys, numpy as np, torch.distributed as dist
sys.path.insert(0, os.environ["PB_ROOT"])
from pybamm_b200 import shard
from tests import util
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
b = util.bundle("spm")
p = util.sample_inputs(b, 9, np.random.default_rng(7))
parts = shard.partition(len(p), world, shard.cost_proxy(p))
mine = parts[rank]
e = util.emu("spm")          # tests only: the emulation stands in for the per-rank GPU solve
r = e.solve(util.T_EVAL, p[mine], t_interp=util.T_INTERP, rtol=1e-6, atol=1e-8, save_vars=True)
payload = (mine, [t[-1] for t in r["t"]], [y[:, 0] for y in r["y"]], r["flag"].tolist())
gathered = [None] * world if rank == 0 else None
dist.gather_object(payload, gathered, dst=0)
if rank == 0:
    full = e.solve(util.T_EVAL, p, t_interp=util.T_INTERP, rtol=1e-6, atol=1e-8, save_vars=True)
    seen = []
    for idx, tend, v, fl in gathered:
        for k, i in enumerate(idx):
            assert tend[k] == full["t"][i][-1] and fl[k] == full["flag"][i]
            np.testing.assert_array_equal(v[k], full["y"][i][:, 0])
            seen.append(int(i))
    assert sorted(seen) == list(range(len(p)))
    print("GLOO_OK", world)
dist.barrier()
dist.destroy_process_group()
'''


def test_sharded_solve_world_size_2_gloo(tmp_path):
    util.emu("spm")   # build once before the ranks race for it
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    env = dict(os.environ, PB_ROOT=ROOT, OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29513", str(script)], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "GLOO_OK 2" in r.stdout


def test_solution_batch_is_a_lazy_list_like_view_of_the_batch_arrays():
    # what solve(inputs=[...]) returns: len / index / negative index / slice / iteration like the reference's list, Solution
    # objects made on access (and kept), all of them views into the arrays of the batch
    from pybamm_b200.solver import IDA_ROOT_RETURN, Solution, SolutionBatch

    n, n_max = 5, 4
    arrays = dict(t=np.arange(n * n_max, dtype=float).reshape(n, n_max), y=np.ones((n, n_max, 1)) * np.arange(n)[:, None, None],
                  yp=None, n_t=np.array([4, 3, 9, 1, 2], np.int32), n_max=n_max, flag=np.array([1, IDA_ROOT_RETURN, 1, -3, IDA_ROOT_RETURN], np.int32),
                  y_term=np.zeros((n, 3)), stats=np.zeros((n, 8), np.int64), event_idx=np.array([-1, 1, -1, -1, 0], np.int32),
                  t_event=np.zeros(n))
    inputs = [{"p": float(i)} for i in range(n)]
    sols = SolutionBatch(arrays, inputs, {"v": slice(0, 1)}, ["low", "high"])
    assert len(sols) == n and not sols._made
    s1 = sols[1]
    assert isinstance(s1, Solution) and sols[1] is s1 and list(sols._made) == [1]
    assert s1.termination == "event: high" and s1.all_inputs == [inputs[1]] and len(s1.t) == 3
    assert np.shares_memory(s1.t, arrays["t"]) and np.shares_memory(s1["v"], arrays["y"])
    assert len(sols[2].t) == n_max                       # more points produced than stored: clipped to the block
    assert sols[-1].termination == "event: low" and sols[3].termination == "failure"
    assert [s.flag for s in sols[1:4]] == [IDA_ROOT_RETURN, 1, -3]
    assert [float(s["v"][0, 0]) for s in sols] == [0.0, 1.0, 2.0, 3.0, 4.0]
    with pytest.raises(IndexError):
        sols[n]
