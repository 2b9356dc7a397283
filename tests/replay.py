"""Replay of the reference's coupled Treynor runs (Test1 / Test2 / Test12 of
examples/Treynor_Iowa_30m/00_Output_Files) around a channels implementation:
met rain series + all-scalar Green-Ampt + EMELI's linear time interpolation
(oracle/coupling.py), everything else Disabled as in the shipped cfg files."""
import os

import numpy as np

from oracle import channels as och
from oracle import coupling as cp
from oracle import d8 as od8

OUTLETS = [(34, 16), (32, 16), (37, 19), (26, 14), (30, 15)]     # June_20_67_outlets.txt (row, col)
CASES = {'kinematic': ('out_Test1_kin_wave', 'Test1_kin_wave', 2.0, 159),
         'diffusive': ('out_Test2_diff_wave', 'Test2_diff_wave', 2.0, 158),
         'dynamic': ('out_Test12_dynam_wave', 'Test12_dynam_wave', 0.5, 140)}
GREEN_AMPT = dict(Ks=7.2e-06, Ki=1e-07, qs=0.485, qi=0.375808, G=0.724)   # June_20_67_infil_green_ampt.cfg


def treynor_grid(golden_dir, name):
    p = os.path.join(golden_dir, 'treynor', name)
    return np.fromfile(p, dtype='>f4').reshape(44, 29).astype(np.float32)


def rain_series(golden_dir):
    p = os.path.join(golden_dir, 'treynor', 'June_20_67_rain_rates.txt')
    return [l.split()[0] for l in open(p).read().splitlines() if l.strip()]


def golden_rows(golden_dir, wave):
    p = os.path.join(golden_dir, 'treynor', CASES[wave][0], 'June_20_67_0D-Q.txt')
    return np.array([l.split() for l in open(p).read().splitlines()[2:]], dtype=float)


def make_oracle(golden_dir, wave):
    b = od8.build(treynor_grid(golden_dir, 'Treynor_DEM.rtg'), 30.0, 30.0)
    inp = och.prepare_inputs(b, treynor_grid(golden_dir, 'Treynor_slope.rtg'),
                             treynor_grid(golden_dir, 'Treynor_chan-w.rtg'),
                             treynor_grid(golden_dir, 'Treynor_chan-a.rtg'),
                             nval=treynor_grid(golden_dir, 'Treynor_chan-n.rtg'))
    wid = {'kinematic': och.KINEMATIC, 'diffusive': och.DIFFUSIVE, 'dynamic': och.DYNAMIC}[wave]
    return och.ChannelsOracle(b, wid, dt=CASES[wave][2], da=np.float64(900.0), slope=inp['slope'],
                              width=inp['width'], angle=inp['angle'], nval=inp['nval'],
                              outlet_ID=OUTLETS[0], n_pixels=1276)


class OracleChannels(cp.Component):
    """The routing oracle behind the framework's calls; records the rows the
    reference writes to <case>_0D-Q.txt (float32, '%15.7f', every 60 s)."""
    provides = ('Q_outlet',)
    uses = (('P_rain', 'met', 'P_rain'), ('IN', 'infil', 'IN'), ('SM', 'snow', 'SM'))

    def __init__(self, ch):
        self.ch, self.dt, self.time = ch, ch.dt, ch.time
        self.time_index, self.status = 0, 'initialized'
        self.Q_outlet = ch.Q_outlet
        self.rows = []

    def update(self):
        ch = self.ch
        ch.P_rain, ch.IN, ch.SM = self.P_rain, self.IN, self.SM
        t = float(ch.time)
        ch.update()
        self.time = ch.time
        if int(t) % 60 == 0:                                   # channels_base.py:2481-2491
            self.rows.append([float('%15.7f' % (t / 60.0))] +
                             [float('%15.7f' % np.float32(ch.Q[o])) for o in OUTLETS])
        self.time_index += 1
        self.status = 'updated'


class BmiChannels(cp.Component):
    """A BMI channels component (the CUDA drop-in) behind the same calls."""
    provides = ('Q_outlet',)
    uses = (('P_rain', 'met', 'P_rain'), ('IN', 'infil', 'IN'), ('SM', 'snow', 'SM'))
    LONG = {'P_rain': 'atmosphere_water__rainfall_volume_flux',
            'IN': 'soil_surface_water__infiltration_volume_flux',
            'SM': 'snowpack__melt_volume_flux'}

    def __init__(self, comp):
        self.comp, self.dt, self.time = comp, comp.dt, comp.time
        self.time_index, self.status = 0, 'initialized'
        self.Q_outlet = comp.get_values('basin_outlet_water_x-section__volume_flow_rate')

    def update(self):
        for attr, long_name in self.LONG.items():
            self.comp.set_values(long_name, getattr(self, attr))
        self.comp.update(-1.0)
        self.time = self.comp.time
        self.time_index += 1
        self.status = self.comp.get_status()


def run(golden_dir, channels, minutes):
    met = cp.MetSeries(60.0, rain_series(golden_dir))
    infil = cp.GreenAmptScalar(60.0, **GREEN_AMPT)
    comps = dict(met=met, chan=channels, infil=infil, snow=cp.Disabled(SM=0))
    fw = cp.Framework(comps, ['met', 'chan', 'snow', 'infil'])     # providers-file order
    while float(channels.time) < minutes * 60 + 4:
        fw.sweep()
    return fw
