"""The provider kernels one by one (csrc/tf_forcing.cu) against the oracle
classes (oracle/providers.py), with every mix of Scalar- and Grid-typed inputs:
same NumPy expressions, same order, so the grids must agree bit for bit."""
import ctypes as C

import numpy as np
import pytest

from oracle import coupling as cp
from oracle import providers as pv

pytestmark = pytest.mark.gpu

NX, NY = 70, 50


class Dev(object):
    """A strip with routing initialised (the providers need the cell areas) and helpers
    to move operands."""

    def __init__(self):
        from topoflow_old_b200 import _lib as L
        from topoflow_old_b200.engine import Strip
        from topoflow_old_b200.routing import setup_channels
        self.L, self.s = L, Strip(NX, NY)
        code = np.zeros((NY, NX), dtype=np.uint8)
        code[1:-1, 1:-1] = 2
        ones = np.ones(NY, dtype=np.float32)
        self.s.d8_set_codes(code, 30 * ones, 30 * ones, 42 * ones)
        setup_channels(self.s, L.KINEMATIC, L.MANNING, 2.0, 900.0, np.full((NY, NX), 0.01), 3.0, 0.0, 0.05)
        self.h = self.s._h

    def grid(self, v=None):
        gid = C.c_int(-1)
        self.L.check(self.L.lib.tf_grid_new(self.h, C.byref(gid)))
        if v is not None:
            if np.ndim(v) == 0:
                self.L.check(self.L.lib.tf_grid_set(self.h, gid.value, None, float(v)))
            else:
                g = np.ascontiguousarray(v, dtype=np.float64)
                self.L.check(self.L.lib.tf_grid_set(self.h, gid.value, self.L.ptr(g, C.c_double), 0.0))
        return gid.value

    def opd(self, v):
        """(gid, scalar) of a 0-d or grid value."""
        return (-1, float(v)) if np.ndim(v) == 0 else (self.grid(v), 0.0)

    def get(self, gid):
        out = np.empty((NY, NX), dtype=np.float64)
        self.L.check(self.L.lib.tf_grid_get(self.h, gid, self.L.ptr(out, C.c_double)))
        return out


@pytest.fixture(scope='module')
def dev():
    d = Dev()
    yield d
    d.s.close()


def _u(seed, lo=0.0, hi=1.0):
    return lo + (hi - lo) * np.random.default_rng(seed).random((NY, NX))


def _as(kind, grid, scalar):
    return grid if kind == 'g' else np.float64(scalar)


def test_met_rain_split(dev):
    for kp, kt in (('g', 'g'), ('g', 's'), ('s', 'g')):
        P, T = _as(kp, _u(1, 0, 3e-5), 2e-5), _as(kt, _u(2, -5, 5), -1.0)
        pr, ps = dev.grid(), dev.grid()
        dev.L.check(dev.L.lib.tf_prov_met_rain(dev.h, *dev.opd(P), *dev.opd(T), pr, ps))
        assert np.array_equal(dev.get(pr), np.broadcast_to(P * (T > 0), (NY, NX)))
        assert np.array_equal(dev.get(ps), np.broadcast_to(P * (T <= 0), (NY, NX)))


@pytest.mark.parametrize('kinds', ['gggg', 'ssss', 'gsgs', 'sgsg'])
def test_snow_degree_day_updates(dev, kinds):
    T = _as(kinds[0], _u(3, -4, 6), 2.5)
    c0 = _as(kinds[1], _u(4, 1, 4), 2.7)
    T0 = _as(kinds[2], _u(5, -1, 1), -0.2)
    rho = _as(kinds[3], _u(6, 200, 400), 300.0)
    h0_snow, h0_swe = _u(7, 0, 0.02), _u(8, 0, 0.006)
    o = pv.SnowDegreeDay(3600.0, c0, T0, h0_snow, h0_swe, rho, np.float64(900.0), NX * NY)
    hs, hw, sm = dev.grid(h0_snow), dev.grid(h0_swe), dev.grid()
    vol_total = 0.0
    for k in range(6):
        Ps = _u(20 + k, 0, 2e-7)
        o.T_air, o.P_snow, o.rho_H2O = (T if np.ndim(T) else cp.scalar(T)), Ps, cp.scalar(1000.0)
        o.update()
        vol = C.c_double()
        dev.L.check(dev.L.lib.tf_prov_snow_degree_day(dev.h, *dev.opd(T), *dev.opd(Ps), *dev.opd(c0), *dev.opd(T0),
                                                      1000.0, *dev.opd(rho), 3600.0, hs, hw, sm, C.byref(vol)))
        vol_total += vol.value
        assert np.array_equal(dev.get(sm), np.broadcast_to(o.SM, (NY, NX))), k
        assert np.array_equal(dev.get(hw), np.broadcast_to(o.h_swe, (NY, NX))), k
        assert np.array_equal(dev.get(hs), np.broadcast_to(o.h_snow, (NY, NX))), k
    assert abs(vol_total - float(o.vol_SM)) <= 1e-12 * float(o.vol_SM) and vol_total > 0


@pytest.mark.parametrize('kinds', ['gggg', 'ssss', 'gsgs'])
def test_priestley_taylor(dev, kinds):
    Qs, Ql = _as(kinds[0], _u(9, 100, 500), 300.0), _as(kinds[1], _u(10, -90, -10), -50.0)
    Ta, Ts = _as(kinds[2], _u(11, -5, 25), 12.0), _as(kinds[3], _u(12, -5, 25), 10.0)
    o = pv.EvapPriestleyTaylor(3600.0, 1.2, _u(13, 0.3, 0.6), 0.05, 0.0, np.float64(900.0), NX * NY, (NY, NX))
    o.Qn_SW, o.Qn_LW, o.T_air, o.T_surf = Qs, Ql, Ta, Ts
    o.update()
    et, vol = dev.grid(), C.c_double()
    dev.L.check(dev.L.lib.tf_prov_evap_priestley_taylor(dev.h, *dev.opd(Qs), *dev.opd(Ql), *dev.opd(Ta), *dev.opd(Ts),
                                                        -1, 1.2, *dev.opd(o.K_soil), -1, 0.05, -1, 0.0, 3600.0, et,
                                                        C.byref(vol)))
    assert np.array_equal(dev.get(et), o.ET)
    assert abs(vol.value - float(o.vol_ET)) <= 1e-12 * float(o.vol_ET) and vol.value > 0


@pytest.mark.parametrize('kinds,water_table', [('ggggg', False), ('sssss', False), ('gsgsg', True)])
def test_green_ampt_updates(dev, kinds, water_table):
    Ks = _as(kinds[0], _u(14, 3e-6, 1e-5), 7.2e-6)
    Ki = _as(kinds[1], _u(15, 5e-8, 2e-7), 1e-7)
    qs = _as(kinds[2], _u(16, 0.45, 0.5), 0.485)
    qi = _as(kinds[3], _u(17, 0.3, 0.4), 0.375808)
    G = _as(kinds[4], _u(18, 0.5, 0.9), 0.724)
    ht = _u(30, 0, 10) if water_table else None
    z = _u(31, 0, 10) if water_table else None
    o = pv.GreenAmpt(60.0, Ks, Ki, qs, qi, G, np.float64(900.0), NX * NY, h_table=ht, elev=z)
    Ig, INg = dev.grid(1e-6), dev.grid()
    gh, gz = (dev.grid(ht), dev.grid(z)) if water_table else (-1, -1)
    vol_total = 0.0
    for k in range(8):
        P = _u(40 + k, 0, 4e-5) * (k % 3 != 2)            # every third update: no rain at all
        SM = _as('g' if k % 2 else 's', _u(60 + k, 0, 2e-6), 1e-6)
        o.P_rain, o.SM = P, (SM if np.ndim(SM) else cp.scalar(SM))
        o.update()
        vol = C.c_double()
        dev.L.check(dev.L.lib.tf_prov_infil_green_ampt(dev.h, *dev.opd(P), *dev.opd(SM), *dev.opd(Ks), *dev.opd(Ki),
                                                       *dev.opd(qs), *dev.opd(qi), *dev.opd(G), gh, gz, 60.0, Ig, INg,
                                                       C.byref(vol)))
        vol_total += vol.value
        assert np.array_equal(dev.get(INg), np.broadcast_to(o.IN, (NY, NX))), k
        assert np.array_equal(dev.get(Ig), np.broadcast_to(o.I, (NY, NX))), k
    assert abs(vol_total - float(o.vol_IN)) <= 1e-12 * float(o.vol_IN)
    if water_table:
        assert (dev.get(INg)[ht >= z] == 0).all() and (ht >= z).any()


def test_emeli_interpolation_rule(dev):
    """a = (v2 - v1) / (t2 - t1), b = v2 - a t2 only when EVERY cell changed
    (time_interpolation.py:91-133); a * t + b as :677."""
    v1, v2 = _u(70), _u(71) + 2.0
    for case in ('linear', 'one cell unchanged', 'same time'):
        x2 = v2.copy()
        t1, t2 = 10.0, 40.0
        if case == 'one cell unchanged':
            x2[17, 23] = v1[17, 23]
        if case == 'same time':
            t2 = t1
        ti = cp.TimeInterpData(v1.copy(), cp.scalar(t1))
        ti.update(x2, cp.scalar(t2))
        g1, g2, ga, gb, out = dev.grid(v1), dev.grid(x2), dev.grid(), dev.grid(), dev.grid()
        flag = C.c_int(-1)
        dev.L.check(dev.L.lib.tf_interp_linear(dev.h, g1, g2, t1, t2, ga, gb, C.byref(flag)))
        assert flag.value == (1 if case == 'linear' else 0)
        assert np.array_equal(dev.get(ga), np.broadcast_to(ti.a, (NY, NX)))
        assert np.array_equal(dev.get(gb), np.broadcast_to(ti.b, (NY, NX)))
        dev.L.check(dev.L.lib.tf_grid_axpb(dev.h, ga, 0.0, gb, 0.0, 23.5, out))
        assert np.array_equal(dev.get(out), np.broadcast_to(ti.value_at(np.float64(23.5)), (NY, NX)))
