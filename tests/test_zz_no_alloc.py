"""GPU tier, collected last: pbgpu_solve_device_noalloc (static device scratch, no allocation in the call) against
pbgpu_solve_device on the same inputs, through the C ABI, on two calls in a row.  Kept in a file of its own at the end of
the collection order so that a failure here cannot hide another test under `pytest -x`.

The DFN image is the one whose solve kernel uses per-team global scratch (BDF history, predictor, Jacobian temporaries);
SPMe (everything in shared memory) covers the queue-only case.
"""
import ctypes as C

import numpy as np
import pytest

from pybamm_b200 import BatchedIDASolver
from pybamm_b200.host import Result
from tests import util

pytestmark = pytest.mark.gpu

STATIC_MAX_SYSTEMS = 1024   # PBGPU_STATIC_MAX_SYSTEMS in include/pbgpu.h


def _device_solve(s, entry, p, te, ti, seed_fill):
    """One call of `entry` (a ctypes function with pbgpu_solve_device's signature) with device buffers."""
    import torch

    dev = torch.device("cuda", 0)
    n = len(p)
    d_te = torch.tensor(te, dtype=torch.float64, device=dev)
    d_ti = torch.tensor(ti, dtype=torch.float64, device=dev)
    d_p = torch.tensor(p, dtype=torch.float64, device=dev)
    n_max = s._lib.pbgpu_result_capacity(len(te), len(ti))
    # poisoned outputs: whatever the call does not write must show up in the comparison
    t_out = torch.full((n, n_max), seed_fill, dtype=torch.float64, device=dev)
    y_out = torch.full((n, n_max, 1), seed_fill, dtype=torch.float64, device=dev)
    n_t = torch.full((n,), -7, dtype=torch.int32, device=dev)
    flag = torch.full((n,), -77, dtype=torch.int32, device=dev)
    res = Result()
    res.n_max, res.n_ret = n_max, 1
    res.t, res.y, res.n_t, res.flag = t_out.data_ptr(), y_out.data_ptr(), n_t.data_ptr(), flag.data_ptr()
    rc = entry(s._solver, n, C.c_void_p(d_te.data_ptr()), len(te), C.c_void_p(d_ti.data_ptr()), len(ti), None, None,
               C.c_void_p(d_p.data_ptr()), 1, C.byref(res), None, 1)
    assert rc == 0, s._lib.pbgpu_last_error().decode(errors="replace")
    assert res.kernel_ms > 0
    torch.cuda.synchronize()
    return n_t.cpu().numpy(), flag.cpu().numpy(), t_out.cpu().numpy(), y_out.cpu().numpy()[:, :, 0]


@pytest.mark.parametrize("name,n", [("dfn", 80), ("spme", 96)])
def test_no_alloc_entry_point_matches_solve_device_twice_in_a_row(name, n):
    b = util.bundle(name)
    s = BatchedIDASolver(b, rtol=1e-8, atol=1e-8, output_variables=[next(iter(b.variables))])
    s._ensure()
    lib = s._lib
    te = np.asarray(util.T_EVAL, dtype=float)
    ti = np.asarray(util.T_INTERP[1:], dtype=float)      # t_eval values removed (canonical form)
    assert lib.pbgpu_solver_set_launch_order(s._solver, None, 0) == 0
    # set-up only: n_sys = 0 launches nothing
    assert lib.pbgpu_solve_device_noalloc(s._solver, 0, None, 0, None, 0, None, None, None, 1, None, None, 0) == 0
    rng = np.random.default_rng(2024)
    for call in range(2):                                 # the second call runs on the first one's leftovers
        p = util.sample_inputs(b, n, rng)
        ref = _device_solve(s, lib.pbgpu_solve_device, p, te, ti, 1e300)
        got = _device_solve(s, lib.pbgpu_solve_device_noalloc, p, te, ti, -1e300)
        np.testing.assert_array_equal(got[0], ref[0])     # number of points
        np.testing.assert_array_equal(got[1], ref[1])     # IDA flags
        assert set(np.unique(ref[1]).tolist()) <= {1, 2}
        for i in range(n):
            k = int(ref[0][i])
            assert k > 10
            np.testing.assert_allclose(got[2][i, :k], ref[2][i, :k], rtol=0, atol=1e-9)    # times [s]
            np.testing.assert_allclose(got[3][i, :k], ref[3][i, :k], rtol=0, atol=1e-9)    # voltage [V]


def test_no_alloc_entry_point_rejects_sizes_beyond_the_static_limit():
    b = util.bundle("spme")
    s = BatchedIDASolver(b, rtol=1e-8, atol=1e-8, output_variables=[next(iter(b.variables))])
    s._ensure()
    res = Result()
    rc = s._lib.pbgpu_solve_device_noalloc(s._solver, STATIC_MAX_SYSTEMS + 1, None, 2, None, 0, None, None, None, 1, C.byref(res),
                                           None, 1)
    assert rc != 0
    msg = s._lib.pbgpu_last_error().decode(errors="replace")
    assert "static limit" in msg and str(STATIC_MAX_SYSTEMS) in msg
