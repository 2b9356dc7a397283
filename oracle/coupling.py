"""Minimal restatement of what surrounds the channels component in a coupled
TopoFlow run (TEST INFRASTRUCTURE ONLY): the EMELI time loop with its linear
time interpolation, the meteorology component's rain time series and the
all-scalar Green-Ampt infiltration.  It exists to replay the reference's
shipped Treynor runs (examples/Treynor_Iowa_30m/00_Output_Files/Test*/) so
that the routing oracle -- and the CUDA component behind the same BMI -- can be
pinned against the reference's own hydrographs.

Follows, including their aliasing quirks:
  framework/emeli.py:1297-1418 (run_model loop), :1841-1933 (get_required_vars)
  framework/time_interpolation.py:36-133 (time_interp_data), :190-293, :468-556,
      :574-685 (initialize / update2 / get_values)
  components/met_base.py:650-700, :1013-1038, :1787-1810 (P series, P_rain)
  components/infil_base.py:216-264, :500-513, :592-647, :700-728, :784-823
  components/infil_green_ampt.py:336-389, Green_Ampt_Infil_Rate_v1
  components/diversions_fraction_method.py:230-252, :418-546, :622-734 (canals)
"""
import numpy as np


def scalar(x, dtype='float64'):
    """A mutable 0-d array, BMI_base.initialize_scalar (utils/BMI_base.py:1941-1972)."""
    return np.array(x, dtype=dtype)


class TimeInterpData(object):
    """time_interpolation.py:36-133.  Note what is a copy and what is a live
    reference: at construction v2 aliases the provider's own object, so a
    variable the provider updates IN PLACE shows its new value in v1 after the
    first update(), while a variable the provider REBINDS keeps its old one."""

    def __init__(self, v1, t1):                                 # time_interpolation.py:63-89
        self.v2 = v1
        self.t2 = t1
        self.v1 = v1.copy()
        self.t1 = t1.copy()
        self.a = np.float64(0)
        self.b = v1

    def update(self, v2, t2):                                   # time_interpolation.py:91-133
        self.t1 = self.t2.copy()
        try:
            self.v1[:] = self.v2.copy()
        except (IndexError, TypeError, ValueError):
            self.v1 = self.v2.copy()
        self.t2 = t2
        try:
            self.v2[:] = v2.copy()
        except (IndexError, TypeError, ValueError):
            self.v2 = v2.copy()
        dv = np.abs(v2 - self.v1)
        if (dv.min() != 0) and (t2 != self.t1):
            self.a = (v2 - self.v1) / (t2 - self.t1)
            self.b = v2 - (self.a * t2)
        else:
            self.a = np.float64(0)
            self.b = v2

    def value_at(self, time):                                   # time_interpolation.py:574-685
        return (self.a * time) + self.b


class Component(object):
    """The slice of BMI_component the loop touches."""
    status = 'initialized'
    provides = ()          # attribute names other components read
    uses = ()              # (attribute, provider component name, provider attribute)

    def __init__(self, dt):                                     # BMI_base.py:1007-1063
        self.dt = scalar(dt)
        self.time = scalar(0)
        self.time_index = 0

    def advance(self):                                          # update_time, BMI_base.py:1067-1114
        self.time_index += 1
        self.time += self.dt
        self.status = 'updated'


class Disabled(Component):
    """A provider switched off in its cfg file: hands out 0-d zeros that never
    change, never advances its clock (its update() returns at once; e.g.
    snow_base.py:193, diversions_base.py:100-126, infil_base.py:223)."""

    def __init__(self, **values):
        Component.__init__(self, 1.0)
        self.provides = tuple(values)
        for k, v in values.items():
            setattr(self, k, scalar(v))

    def update(self):
        pass


class MetSeries(Component):
    """Rain from a one-value-per-line text series [mm/h], dt seconds apart."""
    provides = ('P_rain', 'P_snow', 'T_air', 'rho_H2O')
    MMPH_TO_MPS = np.float64(1) / np.float64(3600000)          # met_base.py:559

    def __init__(self, dt, series_mmph, T_air=20.0, rho_H2O=1000.0):
        Component.__init__(self, dt)
        self.series = [float(v) for v in series_mmph]
        self.next = 0
        self.P = scalar(0)
        self.P_rain = scalar(0)                                 # :608
        self.P_snow = scalar(0)
        self.T_air = scalar(T_air)
        self.rho_H2O = scalar(rho_H2O)
        self.read_input_files()                                 # :626

    def read_input_files(self):                                 # :1787-1810
        if self.next >= len(self.series):
            return                                              # EOF: keep the last value
        # model_input.read_scalar: float32(text); "data *= factor" widens
        data = np.float64(np.float32(self.series[self.next])) * self.MMPH_TO_MPS
        self.next += 1
        self.P[()] = np.float64(data)

    def update(self):                                           # :650-700
        self.P_rain[()] = self.P * (self.T_air > 0)             # :1037-1038, in place
        self.P_snow[()] = self.P * (self.T_air <= 0)            # update_P_snow, in place
        if self.time_index > 0:
            self.read_input_files()
        self.advance()


class MetGrid(Component):
    """Rain from one grid [m/s] that never changes (P_type = Grid)."""
    provides = ('P_rain', 'rho_H2O')

    def __init__(self, dt, P, rho_H2O=1000.0):
        Component.__init__(self, dt)
        self.P = P
        self.P_rain = np.zeros_like(P)
        self.rho_H2O = scalar(rho_H2O)

    def update(self):                                           # met_base.py:650-700, :1037-1038
        self.P_rain[:] = self.P                                 # T_air > 0
        self.advance()


class CanalDiversions(Component):
    """diversions_fraction_method with canals only: what enters a canal leaves
    it nt = ceil(travel_time / dt) updates later.  canals_in_Q comes from the
    channels component; canals_out_Q is updated in place (:701-704)."""
    provides = ('n_canals', 'canals_Q_fraction', 'canals_in_x', 'canals_in_y',
                'canals_out_x', 'canals_out_y', 'canals_out_Q')
    uses = (('canals_in_Q', 'chan', 'canals_in_Q'),)

    def __init__(self, dt, dx, dy, in_cols, in_rows, out_cols, out_rows, fractions, travel_times):
        Component.__init__(self, dt)
        n = len(in_cols)
        self.n_canals = np.array(n, dtype='int32')
        self.canals_Q_fraction = np.array(fractions, dtype='float64')
        self.nt_canals = np.ceil(np.array(travel_times, dtype='float64') / self.dt)   # :509
        self.canals_in_x = np.int32(in_cols) * dx                # :519-522
        self.canals_in_y = np.int32(in_rows) * dy
        self.canals_out_x = np.int32(out_cols) * dx
        self.canals_out_y = np.int32(out_rows) * dy
        self.canal_Q = np.zeros([n, int(self.nt_canals.max())], dtype='float64')      # :533
        self.canals_out_Q = np.zeros([n], dtype='float64')
        self.canals_in_Q = np.zeros([n], dtype='float64')

    def update(self):                                           # :622-734
        self.canal_Q[:, 0] = self.canals_in_Q
        self.canals_out_Q[:] = np.zeros(int(self.n_canals), dtype='float32')
        for k in range(int(self.n_canals)):
            self.canals_out_Q[k] = self.canal_Q[k, int(self.nt_canals[k]) - 1]
        self.canal_Q = np.roll(self.canal_Q, 1, axis=1)
        self.advance()


class SnowDegreeDayScalar(Component):
    """snow_degree_day with every input a scalar: the potential meltrate
    (snow_degree_day.py:272-312), capped by the snow that is there
    (snow_base.py:365-380), then the water-equivalent and depth bookkeeping
    (:397-440).  self.SM is REBOUND at every update, like the reference."""
    provides = ('SM',)
    uses = (('T_air', 'met', 'T_air'), ('P_snow', 'met', 'P_snow'), ('rho_H2O', 'met', 'rho_H2O'))

    def __init__(self, dt, c0, T0, h0_snow, h0_swe, rho_snow, da, n_pixels):
        Component.__init__(self, dt)
        self.c0, self.T0, self.rho_snow = np.float64(c0), np.float64(T0), np.float64(rho_snow)
        self.da, self.n_pixels = np.float64(da), n_pixels
        self.SM = scalar(0)                                     # snow_base.py:323-325
        self.h_snow, self.h_swe = scalar(h0_snow), scalar(h0_swe)
        self.vol_SM = scalar(0)
        self.T_air, self.P_snow, self.rho_H2O = scalar(0), scalar(0), scalar(1000.0)

    def update(self):                                           # snow_base.py:176-233
        M = (self.c0 / np.float64(8.64E7)) * (self.T_air - self.T0)       # [m/s]
        self.SM = np.maximum(M, np.float64(0))
        density_ratio = (self.rho_H2O / self.rho_snow)
        SM_max = (density_ratio / self.dt) * self.h_snow
        self.SM = np.minimum(self.SM, SM_max)
        self.SM = np.maximum(self.SM, np.float64(0))
        volume = np.float64(self.SM * self.da * self.dt)        # update_SM_integral :384-393
        self.vol_SM += (volume * self.n_pixels)
        self.h_swe += (self.P_snow * self.dt)                   # update_swe :419-428
        self.h_swe -= (self.SM * self.dt)
        np.maximum(self.h_swe, np.float64(0), self.h_swe)
        self.h_snow = self.h_swe * density_ratio                # update_depth :432-440
        self.advance()


class GreenAmptScalar(Component):
    """infil_green_ampt with one layer and every input a scalar."""
    provides = ('IN',)
    uses = (('P_rain', 'met', 'P_rain'), ('SM', 'snow', 'SM'))

    def __init__(self, dt, Ks, Ki, qs, qi, G):
        Component.__init__(self, dt)
        self.Ks, self.Ki, self.qs, self.qi, self.G = [np.float64(v) for v in (Ks, Ki, qs, qi, G)]
        self.IN = scalar(0)                                     # infil_green_ampt.py:343-351
        self.I = scalar(1e-6)
        self.vol_IN = scalar(0)
        self.P_rain = scalar(0)
        self.SM = scalar(0)

    def update(self):                                           # infil_base.py:216-264
        self.P_total = (self.P_rain + self.SM)                  # :500-513
        dK = (self.Ks - self.Ki)
        dq = (self.qs - self.qi)
        fc = (self.G * dK * dq / self.I) + self.Ks              # Green_Ampt_Infil_Rate_v1
        self.IN = np.minimum(fc, self.P_total)                  # REBINDS self.IN
        if self.P_total < self.Ks:                              # :784-823 check_low_rainrate
            self.IN = self.P_total
        self.I += (self.IN * self.dt)                           # :700-728
        self.advance()


class Framework(object):
    """emeli.py run_model with time_interp_method = 'Linear'."""

    def __init__(self, comps, order):       # emeli.py:1736-1828, time_interpolation.py:190-293
        self.comps = comps                  # name -> component
        self.order = order                  # provider file order
        self.time = np.float64(0)
        self.dt = min(float(c.dt) for n, c in comps.items() if not isinstance(c, Disabled))
        self.interp = {}
        # initialize_comp_set: hand every provided object to its users by reference
        for name in order:
            user = comps[name]
            for attr, prov, pattr in user.uses:
                setattr(user, attr, getattr(comps[prov], pattr))
        # time_interpolator.initialize(): snapshot, one update, snapshot
        for name in order:
            c = comps[name]
            t1 = c.time.copy()
            for attr in c.provides:
                self.interp[(name, attr)] = TimeInterpData(getattr(c, attr), t1)
            c.update()
            t2 = c.time.copy()
            for attr in c.provides:
                self.interp[(name, attr)].update(getattr(c, attr), t2)

    def _value(self, prov, pattr, time):
        p = self.comps[prov]
        if p.status in ('initialized', 'finalized'):            # time_interpolation.py:590-600
            return getattr(p, pattr)
        return self.interp[(prov, pattr)].value_at(time)

    def sweep(self):
        """One pass of the while-loop body (emeli.py:1359-1418)."""
        for name in self.order:
            c = self.comps[name]
            bmi_time = c.time.copy()
            if self.time > bmi_time:
                for attr, prov, pattr in c.uses:
                    setattr(c, attr, self._value(prov, pattr, bmi_time))
                c.update()
                if c.status != 'initialized':                   # update2, :468-556
                    t2 = c.time.copy()
                    for attr in c.provides:
                        self.interp[(name, attr)].update(getattr(c, attr), t2)
        self.time += self.dt
