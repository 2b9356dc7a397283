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
# fixture directory, name in log_scalars.json, channel dt [s], minute of the last golden row
CASES = {'kinematic': ('out_Test1_kin_wave', 'Test1_kin_wave', 2.0, 158),
         'diffusive': ('out_Test2_diff_wave', 'Test2_diff_wave', 2.0, 157),
         'diffusive_dt6': ('out_Test2_diff_wave_dt6s', 'Test2_diff_wave_dt=6s', 6.0, 157),
         'dynamic': ('out_Test12_dynam_wave', 'Test12_dynam_wave', 0.5, 140),
         # kinematic wave + snow_degree_day as the snow provider (everything else as Test1)
         'kinematic_snow': ('out_Test8_snow_degree_day', 'Test8_snow_degree_day', 2.0, 158)}


def wave_of(case):
    return case.split('_')[0]


def golden_stack(golden_dir, case):
    """<case>_2D-Q.rts of a shipped run: float32 MSB frames of 44 x 29, one per minute."""
    p = os.path.join(golden_dir, 'treynor', CASES[case][0], 'June_20_67_2D-Q.rts')
    return np.fromfile(p, dtype='>f4').reshape(-1, 44, 29).astype(np.float32)
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
    wid = {'kinematic': och.KINEMATIC, 'diffusive': och.DIFFUSIVE, 'dynamic': och.DYNAMIC}[wave_of(wave)]
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
        self.rows, self.frames = [], []

    def update(self):
        ch = self.ch
        ch.P_rain, ch.IN, ch.SM = self.P_rain, self.IN, self.SM
        t = float(ch.time)
        ch.update()
        self.time = ch.time
        if int(t) % 60 == 0:                                   # channels_base.py:2481-2491
            self.rows.append([float('%15.7f' % (t / 60.0))] +
                             [float('%15.7f' % np.float32(ch.Q[o])) for o in OUTLETS])
            self.frames.append(np.float32(ch.Q))               # save_grid_dt = 60 s too
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


def run_to_the_logged_stop(fw, channels, case, log):
    """Carry a replay on to the update at which the reference's driver stopped
    (topoflow_driver, dt = 6 s, n_steps in the console log): the channels
    component has then made (6 / dt) * (n_steps - 1) + 1 updates."""
    dt = CASES[case][2]
    n_updates = int(round(6.0 / dt)) * (int(log['n_steps']) - 1) + 1
    assert channels.time_index <= n_updates
    while channels.time_index < n_updates:
        fw.sweep()
    return n_updates


def cfg_values(path):
    """name -> float for the numeric records of a TopoFlow cfg file."""
    out = {}
    for line in open(path):
        f = [x.strip() for x in line.split('|')]
        if len(f) == 4 and f[2] in ('float', 'double', 'long', 'int'):
            out[f[0]] = float(f[1])
    return out


def degree_day_snow(golden_dir):
    """The snow provider of Test8 (June_20_67_snow_degree_day.cfg; the shipped run only flips
    its comp_status to Enabled)."""
    v = cfg_values(os.path.join(golden_dir, 'treynor', 'June_20_67_snow_degree_day.cfg'))
    return cp.SnowDegreeDayScalar(v['dt'], c0=v['c0'], T0=v['T0'], h0_snow=v['h0_snow'],
                                  h0_swe=v['h0_swe'], rho_snow=v['rho_snow'], da=900.0, n_pixels=1276)


def run(golden_dir, channels, minutes, snow=None):
    met = cp.MetSeries(60.0, rain_series(golden_dir))
    infil = cp.GreenAmptScalar(60.0, **GREEN_AMPT)
    comps = dict(met=met, chan=channels, infil=infil, snow=snow or cp.Disabled(SM=0))
    fw = cp.Framework(comps, ['met', 'chan', 'snow', 'infil'])     # providers-file order
    while float(channels.time) < minutes * 60 + 8:            # just past the last golden row
        fw.sweep()
    return fw


# ---- Test_Plane_Canal/Case1: constant rain grid + one canal, everything else Disabled ----
PLANE_PTS = [(10, 18), (1, 18), (10, 16), (1, 16), (10, 3), (9, 3), (8, 3), (1, 3),
             (3, 5), (2, 5), (1, 5), (1, 1)]                     # Case1_outlets.txt as (row, col)
PLANE_DIV_LONG = {'n_canals': 'canals__count',
                  'canals_Q_fraction': 'canals_entrance_water__volume_fraction',
                  'canals_in_x': 'canals_entrance__x_coordinate',
                  'canals_in_y': 'canals_entrance__y_coordinate',
                  'canals_out_x': 'canals_exit__x_coordinate',
                  'canals_out_y': 'canals_exit__y_coordinate',
                  'canals_out_Q': 'canals_exit_water__volume_flow_rate'}


def plane_grid(golden_dir, name):
    p = os.path.join(golden_dir, 'plane_canal', name)
    return np.fromfile(p, dtype='<f4').reshape(20, 20)


def plane_rain(golden_dir):
    # met_base.py:1799-1800 -> model_input.read_next(..., factor=mmph_to_mps): float32 data
    # scaled in place (float32 arithmetic under NumPy-1 casting), then widened
    rain = plane_grid(golden_dir, 'Rain_100mmph_on_right.rtg')
    return np.float64(rain * np.float32(np.float64(1) / np.float64(3600000)))


def plane_diversions(golden_dir):
    """plane_canals.txt through read_canal_data (diversions_fraction_method.py:418-546)."""
    kv = {}
    for line in open(os.path.join(golden_dir, 'plane_canal', 'plane_canals.txt')):
        if ':' in line:
            k, v = line.split(':', 1)
            kv[k.strip()] = float(v.split()[0])
    assert kv['n_canals'] == 1
    return cp.CanalDiversions(6.0, 10.0, 10.0, [int(kv['canal_in_col'])], [int(kv['canal_in_row'])],
                              [int(kv['canal_out_col'])], [int(kv['canal_out_row'])],
                              [kv['Q_fraction']], [kv['travel_time']])


class OraclePlaneChannels(cp.Component):
    provides = ('Q_outlet', 'canals_in_Q')
    uses = (('P_rain', 'met', 'P_rain'), ('canals_out_Q', 'div', 'canals_out_Q'))

    def __init__(self, golden_dir, div):
        dem, slope = plane_grid(golden_dir, 'plane_DEM.rtg'), plane_grid(golden_dir, 'plane_slope.rtg')
        b = od8.build(dem, 10.0, 10.0)
        ch = och.ChannelsOracle(b, och.KINEMATIC, dt=6.0, da=np.float64(10.0) * np.float64(10.0),
                                slope=np.float64(slope), width=10.0, angle=0.0, nval=0.3,
                                outlet_ID=PLANE_PTS[0])
        self.ch, self.dt, self.time = ch, ch.dt, ch.time
        self.time_index, self.status = 0, 'initialized'
        self.Q_outlet = ch.Q_outlet
        ids = lambda x, y: (np.int32(y / 10.0), np.int32(x / 10.0))     # channels_base.py:1030-1036
        ch.set_canals(*(ids(div.canals_in_x, div.canals_in_y) + ids(div.canals_out_x, div.canals_out_y)
                        + (div.canals_Q_fraction, div.canals_out_Q)))
        self.canals_in_Q = ch.canals_in_Q = np.zeros(1, dtype='float64')
        self.rows = {v: [] for v in 'Qdu'}

    def update(self):
        ch = self.ch
        ch.P_rain, ch.canals_out_Q = self.P_rain, self.canals_out_Q
        t = float(ch.time)
        ch.update()
        self.time, self.canals_in_Q = ch.time, ch.canals_in_Q
        if int(t) % 30 == 0:                                   # channels_base.py:2481-2491
            for v in 'Qdu':                                    # text_ts_files.py:149-186
                vals = np.float32(getattr(ch, v)[tuple(np.array(PLANE_PTS).T)])
                self.rows[v].append([float('%15.7f' % (t / 60.0))] + [float('%15.7f' % x) for x in vals])
        self.time_index += 1
        self.status = 'updated'


class BmiPlaneChannels(cp.Component):
    """The CUDA component in the same seat: the diversions outputs are set into
    it BEFORE initialize(), as emeli.initialize_comp_set does (emeli.py:1797-1828)."""
    provides = ('Q_outlet', 'canals_in_Q')
    uses = (('P_rain', 'met', 'P_rain'), ('canals_out_Q', 'div', 'canals_out_Q'))

    def __init__(self, comp, cfg_file, div):
        for attr, long_name in PLANE_DIV_LONG.items():
            comp.set_values(long_name, getattr(div, attr))
        for long_name in ('sinks__count', 'sources__count'):
            comp.set_values(long_name, np.array(0, dtype='int32'))
        for long_name in ('snowpack__melt_volume_flux', 'land_surface_water__baseflow_volume_flux',
                          'glacier_ice__melt_volume_flux', 'land_surface_water__evaporation_volume_flux',
                          'soil_surface_water__infiltration_volume_flux'):
            comp.set_values(long_name, cp.scalar(0))           # Disabled providers: 0-d zeros
        comp.set_values('water-liquid__mass-per-volume_density', cp.scalar(1000.0))
        comp.initialize(cfg_file=cfg_file, SILENT=True)
        self.comp, self.dt, self.time = comp, comp.dt, comp.time
        self.time_index, self.status = 0, 'initialized'
        self.Q_outlet = comp.get_values('basin_outlet_water_x-section__volume_flow_rate')
        self.canals_in_Q = comp.get_values('canals_entrance_water__volume_flow_rate')

    def update(self):
        c = self.comp
        c.set_values('atmosphere_water__rainfall_volume_flux', self.P_rain)
        c.set_values('canals_exit_water__volume_flow_rate', self.canals_out_Q)
        c.update(-1.0)
        self.time = c.time
        self.canals_in_Q = c.get_values('canals_entrance_water__volume_flow_rate')
        self.time_index += 1
        self.status = c.get_status()


def run_plane(golden_dir, make_channels, seconds=3600):
    div = plane_diversions(golden_dir)
    chan = make_channels(div)
    met = cp.MetGrid(6000.0, plane_rain(golden_dir))
    fw = cp.Framework(dict(met=met, div=div, chan=chan), ['met', 'div', 'chan'])   # Case1_providers.txt
    while float(chan.time) < seconds:
        fw.sweep()
    return fw, chan
