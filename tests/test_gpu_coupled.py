"""A coupled run -- meteorology, degree-day snow, Priestley-Taylor evaporation,
Green-Ampt infiltration and channel routing under EMELI's time loop -- with the
providers on the device (topoflow_old_b200/coupled.py, csrc/tf_forcing.cu)
against the same run of the oracle (oracle/coupling.py's Framework with the
grid providers of oracle/providers.py)."""
import numpy as np
import pytest

from oracle import coupling as cp
from oracle import providers as pv
from tests.cases import make_case, oracle_for
from tests.test_gpu_channels import RTOL, field_err, gpu_for

pytestmark = pytest.mark.gpu


class OracleChan(cp.Component):
    provides = ('Q_outlet',)
    uses = (('P_rain', 'met', 'P_rain'), ('SM', 'snow', 'SM'), ('ET', 'evap', 'ET'), ('IN', 'infil', 'IN'))

    def __init__(self, ch):
        self.ch, self.dt, self.time = ch, ch.dt, ch.time
        self.time_index, self.status = 0, 'initialized'
        self.Q_outlet = ch.Q_outlet
        self.hyd = []

    def update(self):
        ch = self.ch
        ch.P_rain, ch.SM, ch.ET, ch.IN = self.P_rain, self.SM, self.ET, self.IN
        ch.update()
        self.hyd.append([float(ch.Q_outlet), float(ch.u_outlet), float(ch.d_outlet), float(ch.f_outlet)])
        self.time = ch.time
        self.time_index += 1
        self.status = 'updated'


def _forcing(c):
    from topoflow_old_b200.utils import synth
    nx, ny = c['nx'], c['ny']
    u = lambda seed: synth.uniform_field(nx, ny, 0.0, 1.0, seed=seed, dtype=np.float64)
    dem = np.float64(c['dem'])
    dem = np.where(dem < -1000, dem.max(), dem)
    T_air = 6.0 - 12.0 * (dem - dem.min()) / (dem.max() - dem.min())          # +6 .. -6 deg C with elevation
    met = dict(dt=30.0, P=(200.0 / 3.6e6) * (0.5 + u(51)), T_air=T_air, T_surf=T_air - 1.0, Qn_SW=300.0,
               Qn_LW=-50.0, rain_duration=200.0)
    snow = dict(dt=12.0, c0=2.7, T0=-0.2, h0_snow=0.3 * u(52), h0_swe=0.1 * u(52), rho_snow=300.0)
    evap = dict(dt=20.0, alpha=1.2, K_soil=0.45, soil_x=0.05, T_soil_x=0.0)
    infil = dict(dt=6.0, Ks=7.2e-6 * (0.5 + u(53)), Ki=1e-7, qs=0.485, qi=0.375808, G=0.724)
    return met, snow, evap, infil


@pytest.mark.parametrize('wave,dt', [(0, 2.0), (2, 0.5)])
def test_coupled_run_matches_the_oracle_framework(wave, dt):
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200.coupled import CoupledRun
    c = make_case(96, 80, seed=41)
    met, snow, evap, infil = _forcing(c)
    seconds = 300.0 if wave == 0 else 90.0
    n_pix = c['nx'] * c['ny']
    # ---- oracle ----
    ch = oracle_for(c, wave, dt)
    comps = dict(met=pv.MetFields(met['dt'], met['P'], met['T_air'], met['T_surf'], met['Qn_SW'], met['Qn_LW'],
                                  duration=met['rain_duration']),
                 snow=pv.SnowDegreeDay(snow['dt'], snow['c0'], snow['T0'], snow['h0_snow'], snow['h0_swe'],
                                       snow['rho_snow'], c['da'], n_pix),
                 evap=pv.EvapPriestleyTaylor(evap['dt'], evap['alpha'], evap['K_soil'], evap['soil_x'],
                                             evap['T_soil_x'], c['da'], n_pix, (c['ny'], c['nx'])),
                 infil=pv.GreenAmpt(infil['dt'], infil['Ks'], infil['Ki'], infil['qs'], infil['qi'], infil['G'],
                                    c['da'], n_pix),
                 chan=OracleChan(ch))
    fw = cp.Framework(comps, ['met', 'snow', 'evap', 'infil', 'chan'])
    while float(comps['chan'].time) < seconds:
        fw.sweep()
    # ---- device ----
    s = gpu_for(c, wave, dt)
    try:
        run = CoupledRun(s, dt, met, snow, evap, infil)
        run.run_until(seconds)
        assert run.bad[0] == -1 and run.bad[1] == -1
        assert run.clock['chan'][2] == comps['chan'].time_index
        for name in ('snow', 'evap', 'infil', 'met'):
            assert run.clock[name][2] == comps[name].time_index, name
        # provider state and outputs: same expressions, same order
        for got, want, nm in ((run.get(run.SM.obj), comps['snow'].SM, 'SM'), (run.get(run.h_swe), comps['snow'].h_swe, 'h_swe'),
                              (run.get(run.ET.obj), comps['evap'].ET, 'ET'), (run.get(run.IN.obj), comps['infil'].IN, 'IN'),
                              (run.get(run.I), comps['infil'].I, 'I'), (run.get(run.P_rain.obj), comps['met'].P_rain, 'P_rain')):
            assert field_err(got, np.broadcast_to(want, got.shape)) <= 1e-13, nm
        for k, o in (('SM', comps['snow'].vol_SM), ('ET', comps['evap'].vol_ET), ('IN', comps['infil'].vol_IN),
                     ('R', ch.vol_R)):
            assert abs(run.vol[k] - float(o)) <= 1e-11 * max(abs(float(o)), 1e-30), k
        assert float(comps['snow'].vol_SM) > 0 and float(comps['evap'].vol_ET) > 0 and float(comps['infil'].vol_IN) > 0
        # routed state and hydrograph
        for name, fid in (('vol', L.F_VOL), ('d', L.F_D), ('u', L.F_U), ('Q', L.F_Q)):
            assert field_err(s.get(fid), getattr(ch, name)) <= RTOL, name
        hyd = run.hydrograph()[:, 0, :]
        want = np.array(comps['chan'].hyd)
        assert hyd.shape == want.shape
        for j in range(4):
            assert field_err(hyd[:, j], want[:, j]) <= RTOL, j
        assert ch.Q.max() > 0
    finally:
        s.close()
