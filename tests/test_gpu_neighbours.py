"""Pit filling and the saturated-zone lateral-flow gather on the device
(csrc/tf_fill.cu) against their restatements (oracle/neighbours.py)."""
import numpy as np
import pytest

from oracle import d8 as od8
from oracle import neighbours as nb

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('nx,ny,seed,holes', [(60, 50, 3, False), (131, 97, 8, True), (300, 200, 5, False)])
def test_fill_pits_equals_the_priority_flood(nx, ny, seed, holes):
    from topoflow_old_b200.engine import Strip
    from topoflow_old_b200.utils import synth
    dem = synth.fractal_dem(nx, ny, seed=seed, roughness=8.0)
    rng = np.random.default_rng(seed)
    # a few deep artificial pits, one of them several cells wide
    for _ in range(12):
        y, x = rng.integers(2, ny - 2), rng.integers(2, nx - 2)
        dem[y, x] -= np.float32(40.0)
    dem[ny // 2:ny // 2 + 5, nx // 3:nx // 3 + 7] -= np.float32(25.0)
    if holes:
        dem[ny // 4:ny // 4 + 3, nx // 2:nx // 2 + 4] = np.float32(-9999.0)     # a nodata lake: drains its shore
        dem[3 * ny // 4, nx // 5] = np.float32(np.nan)                          # becomes nodata
    want, n_raised = nb.fill_pits(dem)
    with Strip(nx, ny) as s:
        got, n_got, sweeps = s.fill_pits(dem)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert n_got == int((want > np.where(np.isfinite(dem), dem, np.float32(-9999.0))).sum())
    assert n_raised >= n_got > 12 and sweeps >= 8


def test_filled_dem_has_no_pits_for_the_d8_builder():
    from topoflow_old_b200.engine import Strip
    from topoflow_old_b200.utils import synth
    nx, ny = 400, 300
    dem = synth.fractal_dem(nx, ny, seed=12, roughness=8.0)
    ones = np.ones(ny, dtype=np.float32)
    with Strip(nx, ny) as s:
        before = s.d8_build(dem, 30 * ones, 30 * ones, np.float32(np.sqrt(1800.0)) * ones)
        filled, n_raised, _ = s.fill_pits(dem)
        after = s.d8_build(filled, 30 * ones, 30 * ones, np.float32(np.sqrt(1800.0)) * ones)
    assert before['n_pits'] > 0 and n_raised > 0
    # every interior cell now has a neighbour that is not higher: what the D8 builder still calls
    # a pit lies on the grid edge (edge cells are where the water leaves), the rest are flats
    z = filled
    lowest = np.full((ny - 2, nx - 2), np.inf, dtype=np.float32)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            if dy or dx:
                lowest = np.minimum(lowest, z[1 + dy:ny - 1 + dy, 1 + dx:nx - 1 + dx])
    assert (lowest <= z[1:-1, 1:-1]).all()
    assert after['n_pits'] < before['n_pits'] and after['n_pits'] <= 2 * (nx + ny)


@pytest.mark.parametrize('da_kind', ['scalar', 'rows'])
def test_satzone_lateral_flow_gather(da_kind):
    from topoflow_old_b200.engine import Strip
    from topoflow_old_b200.utils import synth
    nx, ny = 120, 90
    dem = synth.fractal_dem(nx, ny, seed=4, roughness=8.0)
    b = od8.build(dem, 30.0, 30.0)
    Q = synth.uniform_field(nx, ny, 0.0, 2e-3, seed=5, dtype=np.float64)
    Rg = synth.uniform_field(nx, ny, 0.0, 1e-6, seed=6, dtype=np.float64)
    if da_kind == 'scalar':
        da, da_arg = np.float64(900.0), 900.0
    else:
        rows = 900.0 * (1.0 + 0.002 * np.arange(ny))
        da, da_arg = np.repeat(rows, nx).reshape(ny, nx), rows
    ones = np.ones(ny, dtype=np.float32)
    with Strip(nx, ny) as s:
        s.d8_build(dem, 30 * ones, 30 * ones, np.float32(np.sqrt(1800.0)) * ones)
        for rg in (Rg, 3e-7):
            want = nb.satzone_dzw(b, Q, rg if np.ndim(rg) else np.float64(rg), da, np.float64(3600.0))
            got = s.satzone_dzw(Q, rg, da_arg, 3600.0)
            assert np.array_equal(got, want)
