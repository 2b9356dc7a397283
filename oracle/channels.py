"""NumPy oracle for the per-timestep channel routing (TEST INFRASTRUCTURE ONLY).

Restates topoflow/components/channels_base.py `update()` and the three
`update_velocity()` variants (channels_kinematic_wave.py:83-113,
channels_diffusive_wave.py:83-130, channels_dynamic_wave.py:92-341) with the
same whole-array NumPy operations in the same order and association, so that
on the same NumPy/libm it produces the reference's values.  Arrays in,
arrays out: file formats, cfg parsing and the BMI shell are not here.

Legacy-NumPy semantics made explicit (the reference ran on NumPy 1.x):
  * `d8.ds = sinu * d8.ds` (channels_base.py:882) stayed float32 when `sinu`
    is a 0-d array (value-based casting); NEP 50 would widen it.
  * scalars from a cfg file are 0-d float64 arrays (BMI_base.py:1592-1598).

Parity: pinned against the reference's shipped Test_Plane_Canal hydrographs
(tests/test_oracle_channels.py).
"""
import numpy as np

G = np.float64(9.81)                     # channels_base.py:456-463
AVAL = np.float64(0.476)
KAPPA = np.float64(0.408)
LAW_CONST = np.sqrt(G) / KAPPA
ONE_THIRD = np.float64(1.0) / 3.0
TWO_THIRDS = np.float64(2.0) / 3.0
DEG_TO_RAD = np.pi / 180.0

KINEMATIC, DIFFUSIVE, DYNAMIC = 0, 1, 2


def scalar(x):
    """0-d float64 'mutable scalar' (BMI_base.py:1941-1972)."""
    return np.array(x, dtype='float64')


def prepare_inputs(d8, slope, width, angle_deg, nval=None, z0val=None):
    """read_input_files() (channels_base.py:2142-2200) for grid inputs that
    were read as float32 and widened by model_input.read_next (:127-172):
    zero widths take the cell's dw, bank angles go from degrees to radians.
    Scalars (0-d) pass through untouched except the angle conversion, which
    the cfg reader does not apply to scalars."""
    out = {}
    out['slope'] = np.float64(slope)
    if np.ndim(width) > 0:
        width = np.float64(width)
        w0 = (width == 0)
        width[w0] = d8['dw'][w0]
    out['width'] = width
    if np.ndim(angle_deg) > 0:
        angle = np.float64(angle_deg)
        angle *= DEG_TO_RAD
    else:
        angle = angle_deg            # scalar angle is used as given (:2186-2192)
    out['angle'] = angle
    if nval is not None:
        out['nval'] = np.float64(nval) if np.ndim(nval) > 0 else nval
    if z0val is not None:
        out['z0val'] = np.float64(z0val) if np.ndim(z0val) > 0 else z0val
    return out


class ChannelsOracle(object):
    """State + update() of channels_component for one wave type.

    d8      : dict from oracle.d8.build() (d8_grid, ds, dw, w, p, parent_IDs,
              noflow_IDs); `ds` is copied before the sinuosity scaling.
    slope, width, angle (radians), nval / z0val, sinu, d0 : 0-d or [ny, nx]
              float64, as they stand after read_input_files().
    """

    def __init__(self, d8, wave, dt, da, slope, width, angle, nval=None,
                 z0val=None, sinu=1.0, d0=0.0, manning=True, law_of_wall=False,
                 outlet_ID=(0, 0), rho_H2O=1000.0, n_pixels=None):
        self.wave = wave
        self.KINEMATIC_WAVE = (wave == KINEMATIC)
        self.MANNING = bool(manning)
        self.LAW_OF_WALL = bool(law_of_wall)
        self.d8 = dict(d8)
        code = d8['d8_grid']
        self.ny, self.nx = code.shape
        self.n_pixels = self.nx * self.ny if n_pixels is None else n_pixels
        self.dt = scalar(dt)
        self.da = da
        self.rho_H2O = scalar(rho_H2O)
        self.outlet_ID = outlet_ID
        self.width = width if np.ndim(width) else scalar(width)
        self.angle = angle if np.ndim(angle) else scalar(angle)
        self.nval = None if nval is None else (nval if np.ndim(nval) else scalar(nval))
        self.z0val = None if z0val is None else (z0val if np.ndim(z0val) else scalar(z0val))
        sinu = sinu if np.ndim(sinu) else scalar(sinu)
        d0 = d0 if np.ndim(d0) else scalar(d0)
        slope = slope if np.ndim(slope) else scalar(slope)
        self.sinu, self.d0 = sinu, d0
        # initialize_computed_vars(), channels_base.py:797-984 -----------------
        ds = d8['ds']
        if np.ndim(sinu) == 0:
            ds = np.float32(sinu) * ds          # :882, legacy value-based cast
        else:
            ds = sinu * ds
        self.ds = ds
        self.dw = d8['dw']
        self.slope = slope / sinu               # :887
        if np.ndim(self.slope) == 0:
            self.slope = np.zeros((self.ny, self.nx)) + self.slope
        self.S_bed = self.slope
        self.S_free = self.S_bed.copy()
        z = lambda: np.zeros((self.ny, self.nx), dtype='float64')
        self.u, self.f, self.d = z(), z(), z()
        self.d += d0                            # :903
        self.Q, self.R = z(), z()
        self.tau, self.u_star, self.froude = z(), z(), z()
        self.vol_R, self.vol_Q = scalar(0), scalar(0)
        if self.KINEMATIC_WAVE:
            self.remove_bad_slopes()            # :949-950
        L2 = self.d * np.tan(self.angle)        # :957-960
        self.A_wet = self.d * (self.width + L2)
        self.P_wet = self.width + (np.float64(2) * self.d / np.cos(self.angle))
        self.vol = self.A_wet * self.ds
        self.Rh = z()
        # forcing, as the Disabled providers hand it over: 0-d zeros
        self.P_rain, self.SM, self.GW = scalar(0), scalar(0), scalar(0)
        self.MR, self.ET, self.IN = scalar(0), scalar(0), scalar(0)
        # diversions (initialize_diversion_vars, :988-1051)
        self.n_sources = self.n_sinks = self.n_canals = 0
        self.canals_in_Q = np.zeros(1, dtype='float64')
        # outlet / peak scalars (:1053-1098)
        self.Q_outlet, self.u_outlet = scalar(0), scalar(0)
        self.d_outlet, self.f_outlet = scalar(0), scalar(0)
        self.Q_peak, self.T_peak = scalar(0), scalar(0)
        self.u_peak, self.Tu_peak = scalar(0), scalar(0)
        self.d_peak, self.Td_peak = scalar(0), scalar(0)
        self.time = scalar(0)
        self.time_min = scalar(0)
        self.time_index = 0
        self.n_overbank = 0

    # ------------------------------------------------------------------
    def remove_bad_slopes(self):
        """channels_base.py:2675-2754."""
        s = self.slope
        wz, wn = (s == 0.0), (s < 0.0)
        wi = np.logical_not(np.isfinite(s))
        good = np.invert(np.logical_or(s <= 0.0, wi))
        nbad = int(wz.sum() + wn.sum() + wi.sum())
        self.n_zero_slope = int(wz.sum())
        if nbad == 0 or not good.any():
            return
        s_min = s[good].min()
        self.S_min, self.S_max = s_min, s[good].max()
        s[wz] = s_min
        s[wn] = s_min
        s[wi] = s_min
        self.slope = np.float64(s)
        self.S_bed = self.slope

    def set_sources(self, rows, cols, Q):
        """source_IDs / sources_Q as initialize_diversion_vars builds them from the
        diversions component's x, y (channels_base.py:1018-1023)."""
        self.n_sources = len(rows)
        self.source_IDs = (np.int32(rows), np.int32(cols))
        self.sources_Q = np.float64(Q)

    def set_sinks(self, rows, cols, Q):
        """channels_base.py:1024-1029."""
        self.n_sinks = len(rows)
        self.sink_IDs = (np.int32(rows), np.int32(cols))
        self.sinks_Q = np.float64(Q)

    def set_canals(self, in_rows, in_cols, out_rows, out_cols, fraction, out_Q):
        """channels_base.py:1030-1044; canals_out_Q is the diversions component's
        own array (diversions_fraction_method.py:545, updated in place :701-704)."""
        self.n_canals = len(in_rows)
        self.canal_in_IDs = (np.int32(in_rows), np.int32(in_cols))
        self.canal_out_IDs = (np.int32(out_rows), np.int32(out_cols))
        self.canals_Q_fraction = np.float64(fraction)
        self.canals_out_Q = np.float64(out_Q)

    # ------------------------------------------------------------------
    def update(self):
        """channels_base.py:575-706, in order."""
        self.update_R()
        self.update_R_integral()
        self.update_discharge()
        self.update_diversions()
        self.update_flow_volume()
        self.update_flow_depth()
        self.update_trapezoid_Rh()
        if not self.KINEMATIC_WAVE:
            self.update_free_surface_slope()
        self.update_shear_stress()
        self.update_shear_speed()
        self.update_friction_factor()
        self.update_velocity()
        self.u[self.d8['noflow_IDs']] = np.float64(0)       # :1783-1791
        wg = self.d_is_pos                                  # :1795-1808
        self.froude[wg] = self.u[wg] / np.sqrt(G * self.d[wg])
        self.froude[self.d_is_neg] = np.float64(0)
        self.update_outlet_values()
        self.update_peak_values()
        self.vol_Q += (self.Q_outlet * self.dt)             # :1885
        ok = self.check()
        # update_time (BMI_base.py:1067-1114)
        self.time_index += 1
        self.time += self.dt
        self.time_min = self.time / np.float64(60)
        return ok

    def update_R(self):                                     # :1103-1147
        self.R = (self.P_rain + self.SM + self.GW + self.MR) - (self.ET + self.IN)

    def update_R_integral(self):                            # :1151-1160
        volume = np.double(self.R * self.da * self.dt)
        if np.size(volume) == 1:
            self.vol_R += (volume * self.n_pixels)
        else:
            self.vol_R += np.sum(volume)

    def update_discharge(self):                             # :1164-1188
        self.Q[:] = self.u * self.A_wet

    def update_diversions(self):                            # :1226-1297
        if self.n_sources > 0:
            self.Q[self.source_IDs] += self.sources_Q
            self.vol[self.source_IDs] += (self.sources_Q * self.dt)
        if self.n_sinks > 0:
            self.Q[self.sink_IDs] -= self.sinks_Q
            self.vol[self.sink_IDs] -= (self.sinks_Q * self.dt)
        if self.n_canals > 0:
            cin = self.canals_Q_fraction * self.Q[self.canal_in_IDs]
            self.canals_in_Q = cin
            self.Q[self.canal_in_IDs] -= cin
            self.vol[self.canal_in_IDs] -= (cin * self.dt)
            self.Q[self.canal_out_IDs] += self.canals_out_Q
            self.vol[self.canal_out_IDs] += (self.canals_out_Q * self.dt)

    def update_flow_volume(self):                           # :1308-1375
        dt = self.dt
        self.vol += (self.R * self.da) * dt
        for k in range(8):
            pk = self.d8['p'][k]
            if pk is not None:
                self.vol[pk] += (dt * self.Q[self.d8['w'][k]])
        self.vol -= (self.Q * dt)
        np.maximum(self.vol, 0.0, self.vol)

    def update_flow_depth(self):                            # :1379-1532
        d, width, angle = self.d, self.width, self.angle
        scalar_angles = (np.size(angle) == 1)
        scalar_widths = (np.size(width) == 1)
        top_width = width + (2.0 * d * np.tan(self.angle))  # :1449
        wb = (top_width > self.dw)
        if not scalar_widths:
            self.width[wb] = self.dw[wb]
        if not scalar_angles:
            self.angle[wb] = 0.0
        self.n_overbank = int(np.sum(wb))
        self.overbank = wb
        if scalar_angles:                                   # :1465-1472
            if angle == 0.0:
                d = self.vol / (width * self.ds)
            else:
                denom = 2.0 * np.tan(angle)
                arg = 2.0 * denom * self.vol / self.ds
                arg += width ** (2.0)
                d = (np.sqrt(arg) - width) / denom
        else:                                               # :1477-1486
            w1 = (angle == 0)
            w2 = np.invert(w1)
            a_top = width[w1] * self.ds[w1]
            d[w1] = self.vol[w1] / a_top
            denom = 2.0 * np.tan(angle[w2])
            arg = 2.0 * denom * self.vol[w2] / self.ds[w2]
            arg += width[w2] ** (2.0)
            d[w2] = (np.sqrt(arg) - width[w2]) / denom
        d[self.d8['noflow_IDs']] = 0.0                      # :1511
        np.maximum(d, 0.0, self.d)                          # :1526
        self.d_is_pos = (self.d > 0)
        self.d_is_neg = np.invert(self.d_is_pos)

    def update_trapezoid_Rh(self):                          # :1590-1643
        d, wb = self.d, self.width
        L2 = d * np.tan(self.angle)
        A_wet = d * (wb + L2)
        P_wet = wb + (np.float64(2) * d / np.cos(self.angle))
        P_wet[self.d8['noflow_IDs']] = np.float64(1)
        Rh = (A_wet / P_wet)
        Rh[self.d8['noflow_IDs']] = np.float64(0)
        self.Rh[:] = Rh
        self.A_wet[:] = A_wet
        self.P_wet[:] = P_wet

    def update_free_surface_slope(self):                    # :1536-1543
        delta_d = (self.d - self.d[self.d8['parent_IDs']])
        self.S_free[:] = self.S_bed + (delta_d / self.ds)

    def update_shear_stress(self):                          # :1567-1577
        slope = self.S_bed if self.KINEMATIC_WAVE else np.abs(self.S_free)
        self.tau[:] = self.rho_H2O * G * self.d * slope

    def update_shear_speed(self):                           # :1581-1586
        self.u_star[:] = np.sqrt(self.tau / self.rho_H2O)

    def update_friction_factor(self):                       # :1654-1721
        wg, wb = self.d_is_pos, self.d_is_neg
        if self.MANNING:
            n2 = self.nval ** np.float64(2)
            if np.ndim(n2) == 0:
                self.f[wg] = G * (n2 / (self.d[wg] ** ONE_THIRD))
            else:
                self.f[wg] = G * (n2[wg] / (self.d[wg] ** ONE_THIRD))
            self.f[wb] = np.float64(0)
        if self.LAW_OF_WALL:
            smoothness = (AVAL / self.z0val) * self.d
            np.maximum(smoothness, np.float64(1.1), smoothness)
            self.f[wg] = (KAPPA / np.log(smoothness[wg])) ** np.float64(2)
            self.f[wb] = np.float64(0)

    def manning_formula(self):                              # :2561-2589
        S = self.S_bed if self.KINEMATIC_WAVE else self.S_free
        return (self.Rh ** TWO_THIRDS) * np.sqrt(S) / self.nval

    def law_of_the_wall(self):                              # :2593-2640
        S = self.S_bed if self.KINEMATIC_WAVE else self.S_free
        smoothness = (AVAL / self.z0val) * self.d
        smoothness = np.maximum(smoothness, np.float64(1.1))
        return LAW_CONST * np.sqrt(self.Rh * S) * np.log(smoothness)

    def update_velocity(self):
        if self.wave == KINEMATIC:              # channels_kinematic_wave.py:83-113
            if self.MANNING:
                self.u[:] = self.manning_formula()
            if self.LAW_OF_WALL:
                self.u[:] = self.law_of_the_wall()
        elif self.wave == DIFFUSIVE:            # channels_diffusive_wave.py:83-130
            self.S_free = np.maximum(self.S_free, 0.0)
            if self.MANNING:
                self.u = self.manning_formula()
            if self.LAW_OF_WALL:
                self.u = self.law_of_the_wall()
        else:                                   # channels_dynamic_wave.py:92-341
            vol, u, Q, f = self.vol, self.u, self.Q, self.f
            u[self.d_is_neg] = 0.0
            Vinv = vol.copy()
            wg = (vol > 0)
            wb = np.invert(wg)
            Vinv[wg] = 1 / vol[wg]
            Vinv[wb] = 0.0
            Ap = self.ds * self.P_wet
            Qtrm = (u * Q)
            grav = G * self.S_free * vol
            fric = (f * u * u * Ap)
            dM = (grav - fric - Qtrm) * self.dt
            M = (u * vol) + dM
            for k in range(8):
                pk = self.d8['p'][k]
                if pk is not None:
                    wk = self.d8['w'][k]
                    M[pk] += u[wk] * Q[wk] * self.dt
            M = np.maximum(M, 0.0)
            u2 = M * Vinv
            u2 = np.maximum(u2, np.float64(0))
            self.u[:] = u2

    def update_outlet_values(self):                         # :1812-1832
        o = self.outlet_ID
        self.Q_outlet.fill(self.Q[o])
        self.u_outlet.fill(self.u[o])
        self.d_outlet.fill(self.d[o])
        self.f_outlet.fill(self.f[o])

    def update_peak_values(self):                           # :1851-1863
        if self.Q_outlet > self.Q_peak:
            self.Q_peak.fill(self.Q_outlet)
            self.T_peak.fill(self.time_min)
        if self.u_outlet > self.u_peak:
            self.u_peak.fill(self.u_outlet)
            self.Tu_peak.fill(self.time_min)
        if self.d_outlet > self.d_peak:
            self.d_peak.fill(self.d_outlet)
            self.Td_peak.fill(self.time_min)

    def check(self):
        """check_flow_depth / check_flow_velocity (:1982-2080): returns
        (True, None) or raises like the reference."""
        d, u = self.d, self.u
        if np.logical_or(d < 0.0, np.logical_not(np.isfinite(d))).any():
            raise RuntimeError('Negative depth found.')
        if np.logical_or(u < 0.0, np.logical_not(np.isfinite(u))).any():
            raise RuntimeError('Negative or NaN velocity found.')
        return True

    def mins_and_maxes(self):                               # :1890-1935
        ny, nx = self.ny, self.nx
        out = {}
        for name in ('Q', 'u', 'd'):
            a = getattr(self, name)[1:(ny - 2) + 1, 1:(nx - 2) + 1]
            out[name + '_min'], out[name + '_max'] = a.min(), a.max()
        return out
