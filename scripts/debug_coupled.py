import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import coupling as cp, providers as pv
from tests.cases import make_case, oracle_for
from tests.test_gpu_channels import gpu_for
from tests.test_gpu_coupled import _forcing, OracleChan
from topoflow_old_b200.coupled import CoupledRun
c = make_case(96, 80, seed=41)
met, snow, evap, infil = _forcing(c)
n_pix = c['nx'] * c['ny']
ch = oracle_for(c, 0, 2.0)
comps = dict(met=pv.MetFields(met['dt'], met['P'], met['T_air'], met['T_surf'], met['Qn_SW'], met['Qn_LW'], duration=met['rain_duration']),
             snow=pv.SnowDegreeDay(snow['dt'], snow['c0'], snow['T0'], snow['h0_snow'], snow['h0_swe'], snow['rho_snow'], c['da'], n_pix),
             evap=pv.EvapPriestleyTaylor(evap['dt'], evap['alpha'], evap['K_soil'], evap['soil_x'], evap['T_soil_x'], c['da'], n_pix, (c['ny'], c['nx'])),
             infil=pv.GreenAmpt(infil['dt'], infil['Ks'], infil['Ki'], infil['qs'], infil['qi'], infil['G'], c['da'], n_pix),
             chan=OracleChan(ch))
fw = cp.Framework(comps, ['met', 'snow', 'evap', 'infil', 'chan'])
s = gpu_for(c, 0, 2.0)
run = CoupledRun(s, 2.0, met, snow, evap, infil)
def cmp(tag):
    bad = []
    for nm, got, want in (('P_rain', run.get(run.P_rain.obj), comps['met'].P_rain), ('SM', run.get(run.SM.obj), comps['snow'].SM),
                          ('h_swe', run.get(run.h_swe), comps['snow'].h_swe), ('ET', run.get(run.ET.obj), comps['evap'].ET),
                          ('IN', run.get(run.IN.obj), comps['infil'].IN), ('I', run.get(run.I), comps['infil'].I)):
        want = np.broadcast_to(want, got.shape)
        if not np.array_equal(got, want):
            d = np.abs(got - want)
            i = np.unravel_index(np.argmax(d), d.shape)
            bad.append('%s: %d cells, worst %s got %r want %r' % (nm, (got != want).sum(), i, got[i], want[i]))
    print(tag, 'clocks', {k: v[2] for k, v in run.clock.items()}, {k: v.time_index for k, v in comps.items()},
          'lin', {k[1]: bool(np.ndim(v.a)) for k, v in fw.interp.items() if k[1] in ('SM', 'ET', 'IN')},
          {'SM': run.SM.linear, 'ET': run.ET.linear, 'IN': run.IN.linear}, 'OK' if not bad else bad)
    return not bad
cmp('init')
for k in range(40):
    v0o, v0d = float(ch.vol_R), run.vol['R']
    fw.sweep(); run.sweep(); run._flush()
    print('   dvol_R oracle %.9g device %.9g  sums %r %r' % (float(ch.vol_R) - v0o, run.vol['R'] - v0d, run._sum0, run._sum1))
    if not cmp('sweep %d t=%g' % (k, float(fw.time))):
        break
print('vol device', run.vol)
print('vol oracle', dict(SM=float(comps['snow'].vol_SM), ET=float(comps['evap'].vol_ET), IN=float(comps['infil'].vol_IN), R=float(ch.vol_R)))
