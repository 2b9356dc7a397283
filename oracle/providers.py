"""TEST INFRASTRUCTURE (oracle): the excess-rain providers with grid-typed
variables, for oracle/coupling.py's Framework.

oracle/coupling.py holds the all-scalar providers that are pinned on the
reference's shipped runs (met rain series + Green-Ampt: Test1 / Test2 / Test12;
degree-day snow: Test8).  The classes here evaluate the same NumPy expressions
with 0-d arrays OR [ny, nx] grids, as the reference does when a cfg file types
a variable "Grid"; tests/test_oracle_providers.py checks that with every input
0-d they reproduce the pinned scalar classes bit for bit.  The Priestley-Taylor
evaporation has no shipped run that can be replayed without the meteorology
component's radiation model (Test6 takes Qn_SW / Qn_LW from it): for that one
class parity is UNPINNED -- it is a restatement of the cited lines only.

What is a rebound and what an in-place output matters to EMELI's time
interpolation (time_interpolation.py:63-133, see coupling.TimeInterpData):
met writes its grids in place (BMI_base.update_var, utils/BMI_base.py:2032-2082),
snow / evaporation / infiltration rebind SM, ET, IN at every update.
"""
import numpy as np

from oracle.coupling import Component, scalar


def _var(v):
    """initialize_var: a 0-d array for a scalar, the float64 grid otherwise."""
    return scalar(float(v)) if np.ndim(v) == 0 else np.array(v, dtype='float64')


def _update_var(var, value):
    """BMI_base.update_var (utils/BMI_base.py:2032-2082): in place, both kinds."""
    if np.ndim(var) == 0:
        var[()] = value
    else:
        var[:] = value


class MetFields(Component):
    """Meteorology with P and T_air given (0-d or grids): P_rain = P * (T_air > 0),
    P_snow = P * (T_air <= 0) (met_base.py:1013-1085).  The rain lasts
    `duration` seconds and is zero afterwards (a two-entry input sequence);
    T_air, T_surf, Qn_SW, Qn_LW are handed on as they are."""
    provides = ('P_rain', 'P_snow', 'T_air', 'T_surf', 'Qn_SW', 'Qn_LW', 'rho_H2O')

    def __init__(self, dt, P, T_air, T_surf=0.0, Qn_SW=0.0, Qn_LW=0.0, duration=np.inf, rho_H2O=1000.0):
        Component.__init__(self, dt)
        self.P0, self.duration = _var(P), float(duration)
        self.P = _var(P)
        self.T_air, self.T_surf = _var(T_air), _var(T_surf)
        self.Qn_SW, self.Qn_LW = _var(Qn_SW), _var(Qn_LW)
        self.rho_H2O = scalar(rho_H2O)
        grid = np.ndim(P) > 0 or np.ndim(T_air) > 0
        shape = np.broadcast(np.asarray(P), np.asarray(T_air)).shape
        self.P_rain = np.zeros(shape, dtype='float64') if grid else scalar(0)       # met_base.py:608-609
        self.P_snow = np.zeros(shape, dtype='float64') if grid else scalar(0)

    def update(self):                                           # met_base.py:650-700
        _update_var(self.P_rain, self.P * (self.T_air > 0))     # :1037-1038
        _update_var(self.P_snow, self.P * (self.T_air <= 0))    # :1081-1082
        if self.time_index > 0:                                 # read_input_files: the next entry
            on = float(self.time + self.dt) < self.duration
            _update_var(self.P, self.P0 if on else np.float64(0) * self.P0)
        self.advance()


class SnowDegreeDay(Component):
    """snow_degree_day (snow_degree_day.py:272-312, snow_base.py:176-233, 365-440)."""
    provides = ('SM',)
    uses = (('T_air', 'met', 'T_air'), ('P_snow', 'met', 'P_snow'), ('rho_H2O', 'met', 'rho_H2O'))

    def __init__(self, dt, c0, T0, h0_snow, h0_swe, rho_snow, da, n_pixels):
        Component.__init__(self, dt)
        self.c0, self.T0, self.rho_snow = _var(c0), _var(T0), _var(rho_snow)
        self.da, self.n_pixels = da, n_pixels
        self.SM = scalar(0)                                     # snow_base.py:323-325
        self.h_snow, self.h_swe = _var(h0_snow), _var(h0_swe)
        self.vol_SM = scalar(0)
        self.T_air, self.P_snow, self.rho_H2O = scalar(0), scalar(0), scalar(1000.0)

    def update(self):
        M = (self.c0 / np.float64(8.64E7)) * (self.T_air - self.T0)
        self.SM = np.maximum(M, np.float64(0))
        density_ratio = (self.rho_H2O / self.rho_snow)
        SM_max = (density_ratio / self.dt) * self.h_snow
        self.SM = np.minimum(self.SM, SM_max)
        self.SM = np.maximum(self.SM, np.float64(0))
        volume = np.float64(self.SM * self.da * self.dt)
        if np.size(volume) == 1:
            self.vol_SM += (volume * self.n_pixels)
        else:
            self.vol_SM += np.sum(volume)
        self.h_swe = self.h_swe + (self.P_snow * self.dt)       # "+=" with a grid on the right
        self.h_swe = self.h_swe - (self.SM * self.dt)
        self.h_swe = np.maximum(self.h_swe, np.float64(0))
        self.h_snow = self.h_swe * density_ratio
        self.advance()


class EvapPriestleyTaylor(Component):
    """evap_priestley_taylor.update_ET_rate (:290-353) + evap_base.update_ET_integral
    (:300-309).  PARITY UNPINNED (see the module docstring)."""
    provides = ('ET',)
    uses = (('T_air', 'met', 'T_air'), ('T_surf', 'met', 'T_surf'), ('Qn_SW', 'met', 'Qn_SW'),
            ('Qn_LW', 'met', 'Qn_LW'))

    def __init__(self, dt, alpha, K_soil, soil_x, T_soil_x, da, n_pixels, shape):
        Component.__init__(self, dt)
        self.alpha, self.K_soil, self.soil_x, self.T_soil_x = _var(alpha), _var(K_soil), _var(soil_x), _var(T_soil_x)
        self.da, self.n_pixels, self.shape = da, n_pixels, shape
        self.ET = np.zeros(shape, dtype='float64')              # evap_base.py: ET starts as a grid of zeros
        self.vol_ET = scalar(0)
        self.T_air = self.T_surf = self.Qn_SW = self.Qn_LW = scalar(0)

    def update(self):
        Q_net = self.Qn_SW + self.Qn_LW
        Qc = self.K_soil * (self.T_soil_x - self.T_surf) / self.soil_x
        Qet = self.alpha * (np.float64(0.406) + (np.float64(0.011) * self.T_air)) * (Q_net - Qc)
        ET = (Qet / np.float64(2.5E+9))
        self.ET = np.maximum(ET, np.float64(0))
        if np.size(self.ET) == 1:                               # :352-353
            self.ET = self.ET + np.zeros(self.shape, dtype='float64')
        volume = np.double(self.ET * self.da * self.dt)
        self.vol_ET += np.sum(volume)
        self.advance()


class GreenAmpt(Component):
    """infil_green_ampt, one layer (infil_green_ampt.py:516-572; infil_base.py:216-264,
    500-513, 592-647, 700-728, 784-823)."""
    provides = ('IN',)
    uses = (('P_rain', 'met', 'P_rain'), ('SM', 'snow', 'SM'))

    def __init__(self, dt, Ks, Ki, qs, qi, G, da, n_pixels, h_table=None, elev=None):
        Component.__init__(self, dt)
        self.Ks, self.Ki, self.qs, self.qi, self.G = [_var(v) for v in (Ks, Ki, qs, qi, G)]
        self.da, self.n_pixels = da, n_pixels
        self.h_table, self.elev = h_table, elev
        self.IN = scalar(0)                                     # infil_green_ampt.py:343-351
        self.I = scalar(1e-6)
        self.vol_IN = scalar(0)
        self.P_rain, self.SM = scalar(0), scalar(0)

    def update(self):
        self.P_total = (self.P_rain + self.SM)
        dK = (self.Ks - self.Ki)
        dq = (self.qs - self.qi)
        fc = (self.G * dK * dq / self.I) + self.Ks
        self.IN = np.minimum(fc, self.P_total)
        if np.size(self.P_total) == 1 and np.size(self.Ks) == 1:            # check_low_rainrate
            if self.P_total < self.Ks:
                self.IN = self.P_total
        else:
            w = np.where(self.P_total < self.Ks)
            if np.size(w[0]) != 0:
                if np.size(self.P_total) > 1:
                    self.IN[w] = self.P_total[w]
                else:
                    self.IN[w] = self.P_total
        if np.size(self.IN) > 1 and self.h_table is not None and self.elev is not None:
            w = np.where(self.h_table >= self.elev)             # adjust_infil_rate :636-639
            if np.size(w[0]) != 0:
                self.IN[w] = np.float64(0)
        volume = np.double(self.IN * self.da * self.dt)         # update_IN_integral :650-660
        if np.size(volume) == 1:
            self.vol_IN += (volume * self.n_pixels)
        else:
            self.vol_IN += np.sum(volume)
        self.I = self.I + (self.IN * self.dt)                   # update_I :711 ("+=", grid on the right)
        self.advance()
