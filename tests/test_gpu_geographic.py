"""Fixed-angle (lat/lon, RTI pixel_geom = 0) grids: every row has its own
dx, dy, dd and cell area (utils/pixels.py:72-240 of the reference), so

  * the D8 slopes still divide by the distances of GLOBAL row 0
    (d8_global.py:209-211) -- also on a strip that does not hold that row;
  * ds, dw and da vary by row (the marching kernels read them per row);
  * the contributing area sums a different pixel area per row.

Everything is compared with the oracle through the C ABI."""
import types

import numpy as np
import pytest

from oracle import channels as och
from oracle import d8 as od8
from tests.test_gpu_channels import RTOL, field_err

pytestmark = pytest.mark.gpu

NX, NY = 150, 70


def _rti():
    return types.SimpleNamespace(pixel_geom=0, xres=np.float64(1.0), yres=np.float64(1.0),
                                 nrows=NY, ncols=NX, y_north_edge=np.float64(41.5))


def _inputs():
    from topoflow_old_b200.utils import pixels, synth
    rti = _rti()
    dx, dy, dd = pixels.get_sizes_by_row(rti, METERS=True)
    # a strong south-north stretch so that rows differ by more than rounding
    dx = dx * np.linspace(1.0, 1.6, NY)
    dd = np.sqrt(dx ** 2 + dy ** 2)
    dx, dy, dd = np.float32(dx), np.float32(dy), np.float32(dd)      # d8_base.py:565-573
    da = np.float64(dx) * np.float64(dy)                              # per row [m^2]
    dem = synth.fractal_dem(NX, NY, seed=21, cell=30.0, roughness=8.0)
    return dem, dx, dy, dd, da


def _oracle_d8(dem, dx, dy, dd):
    d8, stats = od8.flow_grid(dem, dx[0], dy[0], dd[0], link=True, return_stats=True)
    pid = od8.parent_id_grid(d8)
    b = dict(d8_grid=d8, parent_ID_grid=pid, parent_IDs=od8.parent_ids(pid),
             dw=od8.flow_width_grid(d8, dx, dy, dd), ds=od8.flow_length_grid(d8, dx, dy, dd),
             noflow_IDs=od8.noflow_ids(d8), dx=dx, dy=dy, dd=dd, stats=stats)
    b['w'], b['p'] = od8.flow_from_to(d8, pid)
    return b


def test_geographic_d8_area_and_routing_match_oracle():
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200.engine import Strip
    from topoflow_old_b200.routing import setup_channels
    from topoflow_old_b200.utils import synth
    dem, dx, dy, dd, da = _inputs()
    b = _oracle_d8(dem, dx, dy, dd)
    da_grid = np.repeat(da, NX).reshape(NY, NX)                       # pixels.get_da, pixel_geom 0
    width = np.float64(synth.uniform_field(NX, NY, 1.0, 6.0, seed=22))
    angle = np.float64(synth.class_field(NX, NY, [0.0, 30.0, 45.0], seed=23, block=6)) * (np.pi / 180.0)
    nval = np.float64(synth.uniform_field(NX, NY, 0.03, 0.15, seed=24))
    slope = np.full((NY, NX), 0.01) + 0.002 * nval
    with Strip(NX, NY) as s:
        s.d8_build(dem, dx, dy, dd)
        assert np.array_equal(s.d8_get(L.D8_CODE), b['d8_grid'])
        assert np.array_equal(s.d8_get(L.D8_DS).view(np.uint32), b['ds'].view(np.uint32))
        assert np.array_equal(s.d8_get(L.D8_DW).view(np.uint32), b['dw'].view(np.uint32))
        s.d8_area(da, units_km=True)
        want_A = od8.area_grid_fast(b['d8_grid'], da_grid, 'km^2')
        assert np.array_equal(s.d8_get(L.D8_AREA).view(np.uint32), want_A.view(np.uint32))
        for wave, dt in ((L.KINEMATIC, 2.0), (L.DIFFUSIVE, 2.0), (L.DYNAMIC, 0.5)):
            ch = och.ChannelsOracle(b, wave, dt=dt, da=da_grid, slope=slope.copy(), width=width.copy(),
                                    angle=angle.copy(), nval=nval.copy(), outlet_ID=(NY // 2, NX // 2))
            ch.P_rain = och.scalar(60.0 / 3.6e6)
            setup_channels(s, wave, L.MANNING, dt, da, slope.copy(), width.copy(), angle.copy(),
                           nval.copy(), outlets=[(NY // 2, NX // 2)])
            inc = s.set_R(60.0 / 3.6e6)
            _, _, bad = s.step(120)
            for _ in range(120):
                ch.update()
            assert bad[0] == -1 and bad[1] == -1
            assert abs(120 * inc - float(ch.vol_R)) <= 1e-12 * float(ch.vol_R)
            for name, fid in (('vol', L.F_VOL), ('d', L.F_D), ('u', L.F_U), ('Q', L.F_Q)):
                assert field_err(s.get(fid), getattr(ch, name)) <= RTOL, (wave, name)
            assert ch.Q.max() > 0


def test_strip_without_global_row_0_uses_its_sizes():
    """Two strips in one process (halo DEM rows set by hand): the southern one
    must divide by dx, dy, dd of global row 0, not of its own first row."""
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200.engine import Strip
    dem, dx, dy, dd, _ = _inputs()
    want = od8.flow_grid(dem, dx[0], dy[0], dd[0], link=False)
    cut = 31
    got = []
    naive = None
    for r0, r1 in ((0, cut), (cut, NY)):
        with Strip(NX, r1 - r0) as s:
            s.set_strip(r0, NY, r0 > 0, r1 < NY)
            s.d8_set_halo_dem(dem[r0 - 1] if r0 > 0 else None, dem[r1] if r1 < NY else None)
            s.d8_set_row0_sizes(dx[0], dy[0], dd[0])
            s.d8_build(dem[r0:r1], dx[r0:r1], dy[r0:r1], dd[r0:r1], link_flats=False)
            got.append(s.d8_get(L.D8_CODE))
            assert np.array_equal(s.d8_get(L.D8_DS).view(np.uint32),
                                  od8.flow_length_grid(want, dx, dy, dd)[r0:r1].view(np.uint32))
        if r0 > 0:
            with Strip(NX, r1 - r0) as s:               # the same strip left to its own row 0
                s.set_strip(r0, NY, True, False)
                s.d8_set_halo_dem(dem[r0 - 1], None)
                s.d8_build(dem[r0:r1], dx[r0:r1], dy[r0:r1], dd[r0:r1], link_flats=False)
                naive = s.d8_get(L.D8_CODE)
    assert np.array_equal(np.concatenate(got, axis=0), want)
    # the case is a real one: the strip's own sizes would have changed some codes
    assert not np.array_equal(naive, want[cut:])
