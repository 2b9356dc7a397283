"""tf_sum_R_noalloc: the runoff-volume integral for callers that may not
allocate device memory (stream-capture regions, plug-ins living inside a
host application's memory pool, latency-critical loops).

The test hands the entry point device pointers it obtained elsewhere (torch
tensors stand in for "the host application's pool"), calls it twice in a row
through the C ABI, and checks both results against tf_chan_set_R -- the
existing entry point, which keeps its per-row partial sums in a cudaMalloc'ed
buffer -- and against NumPy's own sum((R*da)*dt).  Both sum rows in the same order, so the tolerance of the other
vol_R checks (1e-12 relative) is met with room to spare; the second call runs
on a different grid so stale partials from the first would show.

The file name sorts last on purpose: if this fails under `pytest -x`, every
other test has already run.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-12


def _reference_sum(nx, ny, R, da, dt):
    """(R*da)*dt summed by the existing entry point, tf_chan_set_R."""
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200.engine import Strip
    from topoflow_old_b200.routing import setup_channels
    with Strip(nx, ny) as s:
        code = np.zeros((ny, nx), dtype=np.uint8)
        code[1:-1, 1:-1] = 2
        ones = np.ones(ny, dtype=np.float32)
        s.d8_set_codes(code, 30 * ones, 30 * ones, 42 * ones)
        setup_channels(s, L.KINEMATIC, L.MANNING, dt, da, np.full((ny, nx), 0.01),
                       3.0, 0.0, 0.05)
        return s.set_R(R)


def _noalloc_sum(R_dev, nx, ny, pitch, da, dt, out_dev, da_rows_dev=None, stream=None):
    import torch
    from topoflow_old_b200 import _lib as L
    L.check(L.lib.tf_sum_R_noalloc(C.c_void_p(R_dev.data_ptr()), nx, ny, pitch,
                                   None if da_rows_dev is None else C.c_void_p(da_rows_dev.data_ptr()),
                                   float(da), float(dt), C.c_void_p(out_dev.data_ptr()),
                                   None if stream is None else C.c_void_p(stream.cuda_stream)))
    torch.cuda.synchronize()
    return float(out_dev.item())


def _numpy_sum(R, da, dt):
    """update_R_integral as the reference writes it (channels_base.py:1151-1160)."""
    return float(np.sum((R * da) * dt))


def test_two_calls_match_the_allocating_entry_point():
    import torch
    rng = np.random.default_rng(42)
    out_dev = torch.zeros(1, dtype=torch.float64, device='cuda')
    # call 1: 300 x 517 grid, dense pitch
    R1 = rng.random((300, 517)) * 1e-5 - 2e-6
    want1 = _reference_sum(517, 300, R1, 900.0, 2.0)
    got1 = _noalloc_sum(torch.from_numpy(R1).cuda(), 517, 300, 517, 900.0, 2.0, out_dev)
    assert abs(got1 - want1) <= RTOL * abs(want1)
    assert abs(got1 - _numpy_sum(R1, 900.0, 2.0)) <= RTOL * abs(want1)       # and NumPy itself
    # call 2, straight after, smaller grid inside a padded buffer (pitch > nx)
    R2 = rng.random((41, 64)) * 3e-5
    buf = torch.full((41, 96), float('nan'), dtype=torch.float64, device='cuda')
    buf[:, :64] = torch.from_numpy(R2).cuda()
    want2 = _reference_sum(64, 41, R2, 25.0, 0.5)
    got2 = _noalloc_sum(buf, 64, 41, 96, 25.0, 0.5, out_dev)
    assert abs(got2 - want2) <= RTOL * abs(want2)
    assert abs(got2 - _numpy_sum(R2, 25.0, 0.5)) <= RTOL * abs(want2)
    assert got1 != got2


def test_per_row_cell_areas_on_a_side_stream():
    """Geographic grids have one cell area per row (da_rows_dev), and a caller
    with its own stream passes it: both branches against NumPy, twice."""
    import torch
    rng = np.random.default_rng(7)
    out_dev = torch.zeros(1, dtype=torch.float64, device='cuda')
    stream = torch.cuda.Stream()
    for ny, nx in ((257, 130), (64, 1000)):
        R = rng.random((ny, nx)) * 1e-5 - 1e-6
        da_rows = 900.0 * np.cos(np.linspace(0.6, 0.7, ny))
        R_dev = torch.from_numpy(R).cuda()
        da_dev = torch.from_numpy(da_rows).cuda()
        torch.cuda.synchronize()
        got = _noalloc_sum(R_dev, nx, ny, nx, -1.0, 3.0, out_dev, da_rows_dev=da_dev, stream=stream)
        want = _numpy_sum(R, da_rows[:, None], 3.0)
        assert abs(got - want) <= RTOL * abs(want)


def test_oversized_grid_is_refused():
    import torch
    from topoflow_old_b200 import _lib as L
    out_dev = torch.zeros(1, dtype=torch.float64, device='cuda')
    R = torch.zeros(8, dtype=torch.float64, device='cuda')
    rc = L.lib.tf_sum_R_noalloc(C.c_void_p(R.data_ptr()), 1, 65537, 1, None, 1.0, 1.0,
                                C.c_void_p(out_dev.data_ptr()), None)
    assert rc != 0
    assert b'65536' in L.lib.tf_last_error()
