"""The grid-capable provider oracle (oracle/providers.py) against the pinned
all-scalar one (oracle/coupling.py, pinned on the shipped Treynor runs): with
every input 0-d the two must agree bit for bit, update after update."""
import numpy as np

from oracle import coupling as cp
from oracle import providers as pv


def test_green_ampt_grid_oracle_equals_the_pinned_scalar_one():
    ga = dict(Ks=7.2e-06, Ki=1e-07, qs=0.485, qi=0.375808, G=0.724)
    a = cp.GreenAmptScalar(60.0, **ga)
    b = pv.GreenAmpt(60.0, da=np.float64(900.0), n_pixels=1276, **ga)
    rng = np.random.default_rng(1)
    for k in range(200):
        p = cp.scalar(rng.random() * 3e-5 * (k % 7 != 0))
        s = cp.scalar(rng.random() * 1e-6)
        a.P_rain, a.SM, b.P_rain, b.SM = p, s, p.copy(), s.copy()
        a.update(); b.update()
        assert np.array_equal(a.IN, b.IN) and np.array_equal(a.I, b.I)
    # and a grid of identical cells gives that scalar in every cell
    g = pv.GreenAmpt(60.0, da=np.float64(900.0), n_pixels=12, Ks=np.full((3, 4), ga['Ks']), Ki=ga['Ki'],
                     qs=ga['qs'], qi=ga['qi'], G=ga['G'])
    c = cp.GreenAmptScalar(60.0, **ga)
    for k in range(50):
        p = cp.scalar(2e-5 * (k % 5 != 0))
        g.P_rain, g.SM, c.P_rain, c.SM = p, cp.scalar(0), p, cp.scalar(0)
        g.update(); c.update()
        assert g.IN.shape == (3, 4) and np.all(g.IN == c.IN) and np.all(g.I == c.I)


def test_snow_grid_oracle_equals_the_pinned_scalar_one():
    kw = dict(c0=2.7, T0=-0.2, h0_snow=0.5, h0_swe=0.15, rho_snow=300.0, da=np.float64(900.0), n_pixels=1276)
    a, b = cp.SnowDegreeDayScalar(3600.0, **kw), pv.SnowDegreeDay(3600.0, **kw)
    for k in range(60):
        T = cp.scalar(-3.0 + 0.3 * k)
        for s in (a, b):
            s.T_air, s.P_snow, s.rho_H2O = T, cp.scalar(1e-7 * (k < 10)), cp.scalar(1000.0)
            s.update()
        assert np.array_equal(a.SM, b.SM) and np.array_equal(a.h_swe, b.h_swe)
        assert np.array_equal(a.h_snow, b.h_snow) and np.array_equal(a.vol_SM, b.vol_SM)
    assert float(a.h_swe) < 0.15 and float(a.vol_SM) > 0


def test_met_fields_rain_and_snow_split():
    P = np.array([[1e-5, 2e-5], [3e-5, 0.0]])
    T = np.array([[1.0, -1.0], [0.0, 5.0]])
    m = pv.MetFields(60.0, P, T, duration=150.0)
    m.update()
    assert np.array_equal(m.P_rain, P * (T > 0)) and np.array_equal(m.P_snow, P * (T <= 0))
    for _ in range(4):
        m.update()
    assert not m.P_rain.any() and not m.P_snow.any()           # the rain has ended
