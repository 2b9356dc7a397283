"""Channel-flow component on the GPU: same BMI surface as
topoflow/components/channels_base.py (class channels_component, :151-2757).

What the reference does with ~17 whole-grid NumPy passes per time step
(update_discharge ... update_froude_number, channels_base.py:602-642) is one
call into libtopoflow_b200.so here.  The state grids live on the device; the
host arrays other components hold references to are refreshed in place, and
only when somebody reads them (`get_values`, attribute access, a save step).
The 0-d outlet scalars are refreshed every step, as the framework expects.
"""
import os

import numpy as np

from .. import _lib as L
from ..routing import channel_ds, remove_bad_slopes
from ..utils import BMI_base, model_input, model_output
from . import d8_global as d8_base

# (long name, short name, units, role): role i = input, o = output
_VARS = (
    ('atmosphere_water__rainfall_volume_flux', 'P_rain', 'm s-1', 'i'),
    ('glacier_ice__melt_volume_flux', 'MR', 'm s-1', 'i'),
    ('land_surface_water__baseflow_volume_flux', 'GW', 'm s-1', 'i'),
    ('land_surface_water__evaporation_volume_flux', 'ET', 'm s-1', 'i'),
    ('soil_surface_water__infiltration_volume_flux', 'IN', 'm s-1', 'i'),
    ('snowpack__melt_volume_flux', 'SM', 'm s-1', 'i'),
    ('water-liquid__mass-per-volume_density', 'rho_H2O', 'kg m-3', 'i'),
    ('canals__count', 'n_canals', '1', 'i'),
    ('sinks__count', 'n_sinks', '1', 'i'),
    ('sources__count', 'n_sources', '1', 'i'),
    ('canals_entrance__x_coordinate', 'canals_in_x', 'm', 'i'),
    ('canals_entrance__y_coordinate', 'canals_in_y', 'm', 'i'),
    ('canals_entrance_water__volume_fraction', 'canals_Q_fraction', '1', 'i'),
    ('canals_exit__x_coordinate', 'canals_out_x', 'm', 'i'),
    ('canals_exit__y_coordinate', 'canals_out_y', 'm', 'i'),
    ('canals_exit_water__volume_flow_rate', 'canals_out_Q', 'm3 s-1', 'i'),
    ('sinks__x_coordinate', 'sinks_x', 'm', 'i'),
    ('sinks__y_coordinate', 'sinks_y', 'm', 'i'),
    ('sinks_water__volume_flow_rate', 'sinks_Q', 'm3 s-1', 'i'),
    ('sources__x_coordinate', 'sources_x', 'm', 'i'),
    ('sources__y_coordinate', 'sources_y', 'm', 'i'),
    ('sources_water__volume_flow_rate', 'sources_Q', 'm3 s-1', 'i'),
    ('basin_outlet_water_flow__half_of_fanning_friction_factor', 'f_outlet', '1', 'o'),
    ('basin_outlet_water_x-section__mean_depth', 'd_outlet', 'm', 'o'),
    ('basin_outlet_water_x-section__peak_time_of_depth', 'Td_peak', 'min', 'o'),
    ('basin_outlet_water_x-section__peak_time_of_volume_flow_rate', 'T_peak', 'min', 'o'),
    ('basin_outlet_water_x-section__peak_time_of_volume_flux', 'Tu_peak', 'min', 'o'),
    ('basin_outlet_water_x-section__time_integral_of_volume_flow_rate', 'vol_Q', 'm3', 'o'),
    ('basin_outlet_water_x-section__time_max_of_mean_depth', 'd_peak', 'm', 'o'),
    ('basin_outlet_water_x-section__time_max_of_volume_flow_rate', 'Q_peak', 'm3 s-1', 'o'),
    ('basin_outlet_water_x-section__time_max_of_volume_flux', 'u_peak', 'm s-1', 'o'),
    ('basin_outlet_water_x-section__volume_flow_rate', 'Q_outlet', 'm3 s-1', 'o'),
    ('basin_outlet_water_x-section__volume_flux', 'u_outlet', 'm s-1', 'o'),
    ('canals_entrance_water__volume_flow_rate', 'canals_in_Q', 'm3 s-1', 'o'),
    ('channel_bottom_surface__slope', 'S_bed', '1', 'o'),
    ('channel_bottom_water_flow__domain_max_of_log_law_roughness_length', 'z0val_max', 'm', 'o'),
    ('channel_bottom_water_flow__domain_min_of_log_law_roughness_length', 'z0val_min', 'm', 'o'),
    ('channel_bottom_water_flow__log_law_roughness_length', 'z0val', 'm', 'o'),
    ('channel_bottom_water_flow__magnitude_of_shear_stress', 'tau', 'kg m-1 s-2', 'o'),
    ('channel_bottom_water_flow__shear_speed', 'u_star', 'm s-1', 'o'),
    ('channel_centerline__sinuosity', 'sinu', '1', 'o'),
    ('channel_water__volume', 'vol', 'm3', 'o'),
    ('channel_water_flow__froude_number', 'froude', '1', 'o'),
    ('channel_water_flow__half_of_fanning_friction_factor', 'f', '1', 'o'),
    ('channel_water_flow__domain_max_of_manning_n_parameter', 'nval_max', 'm-1/3 s', 'o'),
    ('channel_water_flow__domain_min_of_manning_n_parameter', 'nval_min', 'm-1/3 s', 'o'),
    ('channel_water_flow__manning_n_parameter', 'nval', 'm-1/3 s', 'o'),
    ('channel_water_surface__slope', 'S_free', '1', 'o'),
    ('channel_water_x-section__domain_max_of_mean_depth', 'd_max', 'm', 'o'),
    ('channel_water_x-section__domain_min_of_mean_depth', 'd_min', 'm', 'o'),
    ('channel_water_x-section__domain_max_of_volume_flow_rate', 'Q_max', 'm3 s-1', 'o'),
    ('channel_water_x-section__domain_min_of_volume_flow_rate', 'Q_min', 'm3 s-1', 'o'),
    ('channel_water_x-section__domain_max_of_volume_flux', 'u_max', 'm s-1', 'o'),
    ('channel_water_x-section__domain_min_of_volume_flux', 'u_min', 'm s-1', 'o'),
    ('channel_water_x-section__hydraulic_radius', 'Rh', 'm', 'o'),
    ('channel_water_x-section__initial_mean_depth', 'd0', 'm', 'o'),
    ('channel_water_x-section__mean_depth', 'd', 'm', 'o'),
    ('channel_water_x-section__volume_flow_rate', 'Q', 'm3 s-1', 'o'),
    ('channel_water_x-section__volume_flux', 'u', 'm s-1', 'o'),
    ('channel_water_x-section__wetted_area', 'A_wet', 'm2', 'o'),
    ('channel_water_x-section__wetted_perimeter', 'P_wet', 'm', 'o'),
    ('channel_x-section_trapezoid_bottom__width', 'width', 'm', 'o'),
    ('channel_x-section_trapezoid_side__flare_angle', 'angle', 'rad', 'o'),
    ('land_surface_water__runoff_volume_flux', 'R', 'm s-1', 'o'),
    ('land_surface_water__domain_time_integral_of_runoff_volume_flux', 'vol_R', 'm3', 'o'),
    ('model__time_step', 'dt', 's', 'o'),
    ('model_grid_cell__area', 'da', 'm2', 'o'),
)

# grids that live on the device: attribute name -> field id of tf_chan_get
_DEVICE_FIELDS = {'vol': L.F_VOL, 'd': L.F_D, 'u': L.F_U, 'Q': L.F_Q, 'f': L.F_F,
                  'Rh': L.F_RH, 'A_wet': L.F_A_WET, 'P_wet': L.F_P_WET, 'tau': L.F_TAU,
                  'u_star': L.F_U_STAR, 'froude': L.F_FROUDE, 'S_free': L.F_S_FREE,
                  'width': L.F_WIDTH, 'angle': L.F_MUTATED, 'R': L.F_R}


class _DeviceGrid(object):
    """Attribute whose host array is brought up to date from the device the
    first time it is read after a time step.  The array object is stable, so
    references held by other components stay valid."""

    def __init__(self, name):
        self.name = name

    def __get__(self, obj, objtype=None):
        if obj is None:
            return self
        return obj._fetch(self.name)

    def __set__(self, obj, value):
        obj.__dict__.setdefault('_host', {})[self.name] = value
        obj.__dict__.setdefault('_stale', set()).discard(self.name)


_ALWAYS_UPLOAD = os.environ.get('TF_B200_ALWAYS_UPLOAD', '') == '1'


def _fingerprint(v):
    """Identity and a sampled content stamp of a flux (0-d or grid)."""
    a = np.asarray(v)
    if a.ndim == 0:
        return ('s', float(a))
    flat = a.reshape(-1)
    step = max(1, flat.size // 512)
    return (id(v), a.shape, flat[::step].tobytes(), flat[-1].tobytes())


class channels_component(BMI_base.BMI_component):

    _input_var_names = [v[0] for v in _VARS if v[3] == 'i']
    _output_var_names = [v[0] for v in _VARS if v[3] == 'o']
    _var_name_map = dict((v[0], v[1]) for v in _VARS)
    _var_units_map = dict((v[0], v[2]) for v in _VARS)

    # -- device-backed attributes ---------------------------------------------------
    def _fetch(self, name):
        host = self.__dict__.get('_host', {})
        if name not in host:
            if not self.__dict__.get('_device_ready', False):
                raise AttributeError(name)
            host[name] = np.zeros([self.ny, self.nx], dtype='float64')
            self._stale.add(name)
        arr = host[name]
        if name in self.__dict__.get('_stale', ()) and np.ndim(arr) == 2:
            self._refresh(name, arr)
            self._stale.discard(name)
        return arr

    def _refresh(self, name, arr):
        strip = self.d8.strip
        if name == 'angle':
            fired = strip.get(L.F_MUTATED)
            arr[fired == 1.0] = 0.0                            # channels_base.py:1454
        elif name == 'S_free' and self.KINEMATIC_WAVE:
            pass                                               # never updated (:625)
        else:
            arr[:] = strip.get(_DEVICE_FIELDS[name])

    def _mark_stale(self):
        stale = self.__dict__.setdefault('_stale', set())
        for name, arr in self.__dict__.get('_host', {}).items():
            if np.ndim(arr) == 2:
                stale.add(name)

    # -- constants, wave type ---------------------------------------------------------
    def set_constants(self):                                   # :451-463
        self.g = np.float64(9.81)
        self.aval = np.float64(0.476)
        self.kappa = np.float64(0.408)
        self.law_const = np.sqrt(self.g) / self.kappa
        self.one_third = np.float64(1.0) / 3.0
        self.two_thirds = np.float64(2.0) / 3.0
        self.deg_to_rad = np.pi / 180.0
        self.rad_to_deg = 180.0 / np.pi

    def set_computed_input_vars(self):                         # :726-752
        ext = self.get_attribute('cfg_extension').lower()
        self.KINEMATIC_WAVE = ('kinematic' in ext)
        self.DIFFUSIVE_WAVE = ('diffusive' in ext)
        self.DYNAMIC_WAVE = ('dynamic' in ext)
        self.code_type = 'Grid'
        self.slope_type = 'Grid'
        self.save_grid_dt = np.maximum(self.save_grid_dt, self.dt)
        self.save_pixels_dt = np.maximum(self.save_pixels_dt, self.dt)

    @property
    def wave_id(self):
        return L.KINEMATIC if self.KINEMATIC_WAVE else (L.DIFFUSIVE if self.DIFFUSIVE_WAVE
                                                        else L.DYNAMIC)

    # -- lifecycle ------------------------------------------------------------------
    def initialize(self, cfg_file=None, mode='nondriver', SILENT=False):    # :467-571
        if not SILENT:
            print(' ')
            print('Channels component: Initializing...')
        self.status = 'initializing'
        self.mode = mode
        self.cfg_file = cfg_file
        self.__dict__['_host'] = {}
        self.__dict__['_stale'] = set()
        self.__dict__['_device_ready'] = False
        self.set_constants()
        self.initialize_config_vars()
        self.read_grid_info()
        self.initialize_basin_vars()
        self.initialize_time_vars()
        if self.comp_status == 'Disabled':
            if not SILENT:
                print('Channels component: Disabled in CFG file.')
            self.SAVE_Q_GRIDS = False
            self.SAVE_Q_PIXELS = False
            self.DONE = True
            self.status = 'initialized'
            return
        # variables that come from files start as grids (:524-540)
        for name in ('slope', 'width', 'angle', 'sinu', 'd0'):
            if getattr(self, name + '_type').lower() != 'scalar':
                setattr(self, name, self.initialize_var(getattr(self, name + '_type'), dtype='float64'))
        if self.MANNING and self.nval_type.lower() != 'scalar':
            self.nval = self.initialize_var(self.nval_type, dtype='float64')
        if self.LAW_OF_WALL and self.z0val_type.lower() != 'scalar':
            self.z0val = self.initialize_var(self.z0val_type, dtype='float64')
        print('Calling initialize_d8_vars()...')
        self.initialize_d8_vars()
        self.open_input_files()
        print('Calling read_input_files()...')
        self.read_input_files()
        print('Calling initialize_computed_vars()...')
        self.initialize_computed_vars()
        self.open_output_files()
        self.status = 'initialized'

    def update(self, dt=-1.0):                                 # :575-706
        self.status = 'updating'
        if self.mode == 'driver':
            self.print_time_and_value(self.Q_outlet, 'Q_out', '[m^3/s]')
        self.update_R()
        self.update_R_integral()
        self.update_diversions()
        taps, canals_in, bad = self.update_channel_state()
        self.update_outlet_values(taps)
        if self.n_canals > 0:
            self.canals_in_Q = canals_in[0].copy()             # :1283-1284
        self.update_peak_values()
        self.update_Q_out_integral()
        D_OK = self.check_flow_depth(bad[0])
        U_OK = self.check_flow_velocity(bad[1])
        OK = (D_OK and U_OK)
        self.write_output_files()
        self.update_time(dt)
        if OK:
            self.status = 'updated'
        else:
            self.status = 'failed'
            self.DONE = True

    def update_until(self, time_sec):
        """Advance to `time_sec` [s] in ONE device call when nothing else has to
        happen in between (no output due, no diversions, forcing unchanged):
        the BMI update_until the reference leaves to the framework
        (BMI_base.py; emeli.py drives update() step by step).  Falls back to
        update() per step otherwise.  Peaks, integrals and the outlet scalars
        are updated from the per-step taps exactly as update() would."""
        dt = float(self.dt)
        n = int(round((float(time_sec) - float(self.time)) / dt))
        while n > 0:
            m = self._steps_without_host_work(n)
            if m <= 1:
                self.update(-1.0)
                n -= 1
                if self.status == 'failed':
                    return
                continue
            self.status = 'updating'
            self.update_R()
            R_inc_before = float(self.vol_R)
            self.update_R_integral()
            inc = float(self.vol_R) - R_inc_before
            taps, canals_in, bad = self.d8.strip.step(m)
            self._mark_stale()
            for k in range(m):
                self.update_outlet_values(taps[k:k + 1])
                self.update_peak_values()
                self.update_Q_out_integral()
                if k > 0:
                    self.vol_R += inc
                self.update_time(-1.0)
            OK = self.check_flow_depth(bad[0]) and self.check_flow_velocity(bad[1])
            self.status = 'updated' if OK else 'failed'
            if not OK:
                self.DONE = True
                return
            n -= m

    def _steps_without_host_work(self, n):
        """How many of the next n steps can be queued as one batch."""
        if int(self.n_sources) > 0 or int(self.n_sinks) > 0 or int(self.n_canals) > 0:
            return 1
        dt, t = float(self.dt), float(self.time)
        grids = self.SAVE_Q_GRIDS or self.SAVE_U_GRIDS or self.SAVE_D_GRIDS or self.SAVE_F_GRIDS
        pixels = self.SAVE_Q_PIXELS or self.SAVE_U_PIXELS or self.SAVE_D_PIXELS or self.SAVE_F_PIXELS
        if self.mode == 'driver':
            return 1
        if grids or pixels:
            # write_output_files (:2465-2491) fires at the steps whose start time is a
            # multiple of the save interval: the batch stops before the next such step
            for k in range(n):
                sec = int(t + k * dt)
                if (grids and sec % int(self.save_grid_dt) == 0) or (pixels and sec % int(self.save_pixels_dt) == 0):
                    return max(k, 1)
        return n

    def finalize(self):                                        # :708-722
        if self.comp_status != 'Disabled':
            self.update_mins_and_maxes(REPORT=False)
        self.print_final_report(comp_name='Channels component')
        self.status = 'finalizing'
        if self.comp_status != 'Disabled':
            self.close_input_files()
            self.close_output_files()
            self.d8.finalize()
        self.status = 'finalized'

    # -- D8 ---------------------------------------------------------------------------
    def initialize_d8_vars(self):                              # :756-793
        self.d8 = d8_base.d8_component()
        self.d8.site_prefix = self.site_prefix
        self.d8.case_prefix = self.case_prefix
        self.d8.in_directory = self.in_directory
        self.d8.device = getattr(self, 'device', 0)
        self.d8.initialize(cfg_file=None, SILENT=self.SILENT, REPORT=self.REPORT)
        self.d8.update_most(self.time, SILENT=False, REPORT=True)
        self.d8.update_noflow_IDs()

    # -- initial state -------------------------------------------------------------------
    def initialize_computed_vars(self):                        # :797-984
        if self.MANNING:
            if self.nval is not None:
                self.nval_min = self.nval.min()
                self.nval_max = self.nval.max()
                print('    min(nval) =', self.nval_min)
                print('    max(nval) =', self.nval_max)
            self.z0val = self.initialize_scalar(-1, dtype='float64')
            self.z0val_min = self.initialize_scalar(-1, dtype='float64')
            self.z0val_max = self.initialize_scalar(-1, dtype='float64')
        if self.LAW_OF_WALL:
            if self.z0val is not None:
                self.z0val_min = self.z0val.min()
                self.z0val_max = self.z0val.max()
                print('    min(z0val) =', self.z0val_min)
                print('    max(z0val) =', self.z0val_max)
            self.nval = self.initialize_scalar(-1, dtype='float64')
            self.nval_min = self.initialize_scalar(-1, dtype='float64')
            self.nval_max = self.initialize_scalar(-1, dtype='float64')
        if not self.MANNING and not self.LAW_OF_WALL:
            raise RuntimeError('In CFG file, MANNING=0 and LAW_OF_WALL=0: the reference then '
                               'never updates the velocity (channels_base.py:834-843); '
                               'choose one friction law.')
        width, angle = self._host['width'], self._host['angle']
        print('    min(width)      =', width.min())
        print('    max(width)      =', width.max())
        print('    min(sinuosity)  =', self.sinu.min())
        print('    max(sinuosity)  =', self.sinu.max())
        print('    min(init_depth) =', self.d0.min())
        print('    max(init_depth) =', self.d0.max())
        self.d8.ds = channel_ds(self.sinu, self.d8.ds)         # :882
        self.slope = (self.slope / self.sinu)                  # :887
        self.S_bed = self.slope
        s_free0 = self.S_bed.copy()                            # :889
        print('Initializing u, f, d grids...')
        d = self.initialize_grid(0, dtype='float64')
        d += self.d0                                           # :903
        self.R = self.initialize_grid(0, dtype='float64')
        self.vol_R = self.initialize_scalar(0, dtype='float64')
        self.vol_Q = self.initialize_scalar(0, dtype='float64')
        if self.KINEMATIC_WAVE:
            report = {}
            fixed = remove_bad_slopes(self.slope, report)      # :949-950, :2675-2754
            if report:
                self._print_bad_slope_warning(report)
            self.slope = fixed
            # the reference leaves S_bed pointing at the array it repaired in place
        tan_a, cos_a = np.tan(angle), np.cos(angle)
        L2 = d * tan_a                                         # :957-960
        A_wet = d * (width + L2)
        vol = A_wet * self.d8.ds
        # ---- hand everything to the device ----
        strip = self.d8.strip
        rough = self.nval if self.MANNING else self.z0val
        strip.set_param(L.P_WIDTH, width, is_grid=np.ndim(width) > 0)
        strip.set_param(L.P_TAN_ANGLE, tan_a, is_grid=np.ndim(angle) > 0)
        strip.set_param(L.P_COS_ANGLE, cos_a, is_grid=np.ndim(angle) > 0)
        strip.set_param(L.P_SLOPE, self.S_bed, is_grid=True)
        strip.set_param(L.P_ROUGH, rough, is_grid=np.ndim(rough) > 0)
        strip.set_param(L.P_DS, np.float64(self.d8.ds), is_grid=True)
        outlets = np.stack([self.outlet_IDs[0], self.outlet_IDs[1]], axis=1)
        da = np.asarray(self.da, dtype='float64')
        da_rows = da[:, 0] if da.ndim == 2 else da
        strip.chan_init(self.wave_id, L.MANNING if self.MANNING else L.LAW_OF_WALL,
                        float(self.dt), da_rows, outlets=outlets,
                        allow_packed=getattr(self, 'ALLOW_PACKED', True))
        strip.set_state(L.F_VOL, vol)
        strip.set_state(L.F_D, d)
        # host views: d, vol, A_wet are known; the rest start at zero (:900-967)
        self._host.update(d=d, vol=vol, A_wet=A_wet, S_free=s_free0)
        self._host['P_wet'] = width + (np.float64(2) * d / cos_a)
        for name in ('u', 'f', 'Q', 'tau', 'u_star', 'froude', 'Rh'):
            self._host[name] = self.initialize_grid(0, dtype='float64')
        self.__dict__['_device_ready'] = True
        self._stale.clear()
        self.initialize_diversion_vars()
        self.initialize_outlet_values()
        self.initialize_peak_values()
        self.initialize_min_and_max_values()

    def _print_bad_slope_warning(self, r):                     # :2727-2742
        print('------------------------------------------------------')
        print('WARNING:  Found grid cells with invalid slope.')
        print('  ' + str(r['nz']) + ' grid cells have slope = 0.')
        print('  ' + str(r['nn']) + ' grid cells have slope < 0.')
        print('  ' + str(r['ni']) + ' grid cells have slope of NaN or Inf.')
        print('  For Kinematic Wave routing, these grid cells would')
        print('  have an invalid flow velocity.  Replacing them with')
        print('  smallest positive slope as work-around.')
        print('    min(S) = ' + str(r['S_min']))
        print('    max(S) = ' + str(r['S_max']))
        print('------------------------------------------------------')
        print(' ')

    def initialize_diversion_vars(self):                       # :988-1051
        for name in ('n_sources', 'n_sinks', 'n_canals'):
            if not hasattr(self, name):
                setattr(self, name, 0)
        n = int(self.n_canals)
        self.canals_in_Q = np.zeros(max(n, 1), dtype='float64')
        ids = lambda x, y: (np.int32(np.asarray(y) / self.dy), np.int32(np.asarray(x) / self.dx))
        if int(self.n_sources) > 0:
            self.source_IDs = ids(self.sources_x, self.sources_y)
        if int(self.n_sinks) > 0:
            self.sink_IDs = ids(self.sinks_x, self.sinks_y)
        if n > 0:
            self.canal_in_IDs = ids(self.canals_in_x, self.canals_in_y)
            self.canal_out_IDs = ids(self.canals_out_x, self.canals_out_y)

    def initialize_outlet_values(self):                        # :1053-1068
        for name in ('Q_outlet', 'u_outlet', 'd_outlet', 'f_outlet'):
            setattr(self, name, self.initialize_scalar(0, dtype='float64'))

    def initialize_peak_values(self):                          # :1072-1082
        for name in ('Q_peak', 'T_peak', 'u_peak', 'Tu_peak', 'd_peak', 'Td_peak'):
            setattr(self, name, self.initialize_scalar(0, dtype='float64'))

    def initialize_min_and_max_values(self):                   # :1086-1098
        v = 1e6
        for name in ('Q', 'u', 'd'):
            setattr(self, name + '_min', self.initialize_scalar(v, dtype='float64'))
            setattr(self, name + '_max', self.initialize_scalar(-v, dtype='float64'))

    # -- per-step pieces ------------------------------------------------------------------
    def update_R(self):                                        # :1103-1147
        """All six fluxes 0-d (what Disabled providers and a uniform rain give):
        R is a host scalar and travels as a kernel argument.  Otherwise the
        fluxes go to the device as they are and R is formed there
        (tf_chan_assemble_R) -- no NumPy temporaries, no grid-sized R upload."""
        fluxes = (self.P_rain, self.SM, self.GW, self.MR, self.ET, self.IN)
        if all(np.size(v) == 1 for v in fluxes):
            self.R = (self.P_rain + self.SM + self.GW + self.MR) - (self.ET + self.IN)
            self.__dict__['_R_on_device'] = False
        else:
            # EMELI sets every flux into the component before every update
            # (emeli.py:1841-1933), changed or not.  A flux goes to the device
            # only when it IS a different array or its contents changed: the
            # test is the array's identity plus a fingerprint of 512 samples
            # spread over the grid, which catches what providers do (they
            # rewrite whole grids) at a cost independent of the grid size.
            # TF_B200_ALWAYS_UPLOAD=1 restores the unconditional copy.
            strip = self.d8.strip
            seen = self.__dict__.setdefault('_flux_seen', {})
            changed = False
            for k, v in enumerate(fluxes):                     # L.R_P_RAIN .. L.R_IN
                stamp = _fingerprint(v)
                if seen.get(k) != stamp or _ALWAYS_UPLOAD:
                    strip.set_forcing(k, v)
                    seen[k] = stamp
                    changed = True
            self.__dict__['_R_changed'] = changed or not self.__dict__.get('_R_on_device', False)
            self.__dict__['_R_on_device'] = True

    def update_R_integral(self):                               # :1151-1160
        strip = self.d8.strip
        if not self.__dict__.get('_R_on_device', False):
            strip.set_R(float(self.R), want_sum=False)
            volume = np.double(self.R * self.da * self.dt)
            self.vol_R += (volume * self.rti.n_pixels)
        elif not self.__dict__.get('_R_changed', True):
            self.vol_R += self.__dict__['_R_sum']              # same fluxes, same R, same integral
        else:
            self.__dict__['_R_sum'] = strip.assemble_R(want_sum=True)
            self.vol_R += self.__dict__['_R_sum']
            host = self.__dict__['_host']
            if np.ndim(host.get('R')) != 2:
                host['R'] = np.zeros([self.ny, self.nx], dtype='float64')
            self._stale.add('R')
        rho = getattr(self, 'rho_H2O', None)
        if rho is not None and np.size(rho) == 1:
            strip.set_rho(float(rho))

    def update_diversions(self):                               # :1226-1297
        """The point sets are a few numbers: they are compared with what the
        device already holds and uploaded only when something changed."""
        sets = []
        if int(self.n_sources) > 0:
            sets.append((L.PT_SOURCES, self.source_IDs, self.sources_Q))
        if int(self.n_sinks) > 0:
            sets.append((L.PT_SINKS, self.sink_IDs, self.sinks_Q))
        if int(self.n_canals) > 0:
            sets.append((L.PT_CANALS_IN, self.canal_in_IDs, self.canals_Q_fraction))
            sets.append((L.PT_CANALS_OUT, self.canal_out_IDs, self.canals_out_Q))
        seen = self.__dict__.setdefault('_points_seen', {})
        for kind, ids, vals in sets:
            key = (np.asarray(ids[0]).tobytes(), np.asarray(ids[1]).tobytes(),
                   np.asarray(vals, dtype=np.float64).tobytes())
            if seen.get(kind) != key:
                self.d8.strip.set_points(kind, ids[0], ids[1], vals)
                seen[kind] = key

    def update_channel_state(self):
        """update_discharge, update_flow_volume, update_flow_depth,
        update_trapezoid_Rh, update_free_surface_slope, update_shear_stress /
        speed, update_friction_factor, update_velocity, update_velocity_on_edges
        and update_froude_number (:602-642) as one device step."""
        out = self.d8.strip.step(1)
        self._mark_stale()
        return out

    def update_outlet_values(self, taps):                      # :1812-1832
        self.__dict__['_taps'] = taps[0]
        Q, u, d, f = taps[0, 0]
        self.Q_outlet.fill(Q)
        self.u_outlet.fill(u)
        self.d_outlet.fill(d)
        self.f_outlet.fill(f)

    def update_peak_values(self):                              # :1851-1863
        for val, peak, when in ((self.Q_outlet, self.Q_peak, self.T_peak),
                                (self.u_outlet, self.u_peak, self.Tu_peak),
                                (self.d_outlet, self.d_peak, self.Td_peak)):
            if val > peak:
                peak.fill(val)
                when.fill(self.time_min)

    def update_Q_out_integral(self):                           # :1879-1885
        self.vol_Q += (self.Q_outlet * self.dt)

    def update_mins_and_maxes(self, REPORT=False):             # :1890-1977
        mm = self.d8.strip.minmax()
        for k, name in enumerate(('Q', 'u', 'd')):
            lo, hi = getattr(self, name + '_min'), getattr(self, name + '_max')
            if mm[2 * k] < lo:
                lo.fill(mm[2 * k])
            if mm[2 * k + 1] > hi:
                hi.fill(mm[2 * k + 1])
        if REPORT:
            print('In channels_base.update_mins_and_maxes():')
            print('(dmin, dmax) =', self.d_min, self.d_max)
            print('(umin, umax) =', self.u_min, self.u_max)
            print('(Qmin, Qmax) =', self.Q_min, self.Q_max)
            print(' ')

    def _abort(self, what, label, grid, first_bad):
        bad = np.logical_or(grid < 0.0, np.logical_not(np.isfinite(grid)))
        nbad = int(bad.sum())
        star = '*******************************************'
        for line in (star, 'ERROR: Simulation aborted.', ' ',
                     what + ': ' + str(grid[bad].min() if nbad else 'nan'),
                     'Time step may be too large.', 'Time step:      ' + str(self.dt) + ' [s]'):
            print(line)
        if nbad < 30:
            brow, bcol = divmod(int(first_bad), self.nx)
            print(' ')
            print('(Column, Row):  ' + str(bcol) + ', ' + str(brow))
            print(label + str(grid[brow, bcol]))
        print(star)
        print(' ')

    def check_flow_depth(self, first_bad=-1):                  # :1982-2030
        if first_bad < 0:
            return True
        self._abort('Negative depth found', 'Flow depth:     ', self.d, first_bad)
        raise RuntimeError('Negative depth found.')

    def check_flow_velocity(self, first_bad=-1):               # :2034-2080
        if first_bad < 0:
            return True
        self._abort('Negative or NaN velocity found', 'Velocity:       ', self.u, first_bad)
        raise RuntimeError('Negative or NaN velocity found.')

    # -- input files ----------------------------------------------------------------------
    _IN_FILES = ('slope', 'nval', 'z0val', 'width', 'angle', 'sinu', 'd0')

    def _uses(self, name):
        return (name != 'nval' or self.MANNING) and (name != 'z0val' or self.LAW_OF_WALL)

    def open_input_files(self):                                # :2105-2138
        self.prepend_directory([n + '_file' for n in self._IN_FILES], INPUT=True)
        for name in self._IN_FILES:
            if self._uses(name):
                setattr(self, name + '_unit',
                        model_input.open_file(getattr(self, name + '_type'),
                                              getattr(self, name + '_file')))

    def read_input_files(self):                                # :2142-2200
        rti = self.rti
        for name in self._IN_FILES:
            if not self._uses(name):
                continue
            value = model_input.read_next(getattr(self, name + '_unit'),
                                          getattr(self, name + '_type'), rti)
            if value is None:
                continue
            if name == 'width':
                w0 = (value == 0)                              # zero widths take the flow width
                value[w0] = self.d8.dw[w0]
            elif name == 'angle':
                value *= self.deg_to_rad                       # degrees -> radians
            self.update_var(name, value)

    def close_input_files(self):                               # :2292-2310
        for name in self._IN_FILES:
            unit = getattr(self, name + '_unit', None)
            if unit is not None and self._uses(name) and getattr(self, name + '_type') != 'Scalar':
                unit.close()

    # -- output files -----------------------------------------------------------------------
    _OUT = (('Q', 'volumetric_discharge', 'm^3/s'), ('u', 'mean_channel_flow_velocity', 'm/s'),
            ('d', 'max_channel_flow_depth', 'm'), ('f', 'friction_factor', 'none'))

    def update_outfile_names(self):                            # :2324-2337
        for v, _, _ in self._OUT:
            for kind in ('_gs_file', '_ts_file'):
                setattr(self, v + kind, self.out_directory + getattr(self, v + kind))

    def open_output_files(self):                               # :2378-2461
        model_output.check_netcdf()
        self.update_outfile_names()
        for v, long_name, units in self._OUT:
            if getattr(self, 'SAVE_%s_GRIDS' % v.upper()):
                model_output.open_new_gs_file(self, getattr(self, v + '_gs_file'), self.rti,
                                              var_name=v, long_name=long_name, units_name=units)
            if getattr(self, 'SAVE_%s_PIXELS' % v.upper()):
                model_output.open_new_ts_file(self, getattr(self, v + '_ts_file'), self.outlet_IDs,
                                              var_name=v, long_name=long_name, units_name=units)

    def write_output_files(self, time_seconds=None):           # :2465-2491
        if time_seconds is None:
            time_seconds = self.time_sec
        model_time = int(time_seconds)
        if model_time % int(self.save_grid_dt) == 0:
            self.save_grids()
        if model_time % int(self.save_pixels_dt) == 0:
            self.save_pixel_values()

    def close_output_files(self):                              # :2503-2513
        self.flush_frames()
        self.__dict__.pop('_frame_bufs', None)                 # pinned memory goes with the strip
        for v, _, _ in self._OUT:
            if getattr(self, 'SAVE_%s_GRIDS' % v.upper(), False):
                model_output.close_gs_file(self, v)
            if getattr(self, 'SAVE_%s_PIXELS' % v.upper(), False):
                model_output.close_ts_file(self, v)

    _FRAME_FIELD = {'Q': L.F_Q, 'u': L.F_U, 'd': L.F_D, 'f': L.F_F}

    def save_grids(self):                                      # :2517-2535
        """rts_file.add_grid's float32 cast and byte swap happen on the device
        (tf_chan_frame); the frame comes back on a copy stream into page-locked
        memory while the next steps run, and is appended to the stack at the
        next save (or at close).  The files end up byte-identical."""
        self.flush_frames()
        strip = self.d8.strip
        pend = self.__dict__.setdefault('_frames', [])
        bufs = self.__dict__.setdefault('_frame_bufs', {})
        for v, _, _ in self._OUT:
            if not getattr(self, 'SAVE_%s_GRIDS' % v.upper()):
                continue
            unit = getattr(self, v + '_rts_unit', None)
            if unit is None:
                continue
            if np.dtype(unit.dtype) != np.float32:
                model_output.add_grid(self, getattr(self, v), v, self.time_min)
                continue
            if v not in bufs:
                bufs[v] = strip.pinned_frame()
            strip.frame(self._FRAME_FIELD[v], unit.info.SWAP_ENDIAN, out=bufs[v], wait=False)
            pend.append((unit, bufs[v]))

    def flush_frames(self):
        pend = self.__dict__.get('_frames')
        if pend:
            self.d8.strip.frame_wait()
            for unit, buf in pend:
                unit.add_frame(buf)
            del pend[:]

    def save_pixel_values(self):                               # :2539-2557
        """The monitored cells are exactly the outlet taps the step returned,
        so no grid has to come back from the device for these rows."""
        taps = self.__dict__.get('_taps')
        for j, (v, _, _) in enumerate(self._OUT):
            if getattr(self, 'SAVE_%s_PIXELS' % v.upper()):
                unit = getattr(self, v + '_tts_unit', None)
                if unit is not None and taps is not None:
                    unit.add_values(self.time_min, np.float32(taps[:, j]))

    # the reference's friction laws, kept for callers that use them directly
    def manning_formula(self):                                 # :2561-2589
        S = self.S_bed if self.KINEMATIC_WAVE else self.S_free
        return (self.Rh ** self.two_thirds) * np.sqrt(S) / self.nval

    def law_of_the_wall(self):                                 # :2593-2640
        S = self.S_bed if self.KINEMATIC_WAVE else self.S_free
        smoothness = np.maximum((self.aval / self.z0val) * self.d, np.float64(1.1))
        return self.law_const * np.sqrt(self.Rh * S) * np.log(smoothness)


for _name in _DEVICE_FIELDS:
    setattr(channels_component, _name, _DeviceGrid(_name))
