"""A coupled run with the excess-rain providers on the device.

In the reference, EMELI's time loop (framework/emeli.py:1359-1418) calls every
component's update() in provider order whenever the framework clock has passed
the component's own, hands each user the linearly interpolated outputs of its
providers (framework/time_interpolation.py:63-133, 574-685) as freshly made host
arrays, and so moves four grids into the channels component at every channel
step.  `CoupledRun` runs the same loop -- same order, same clocks, same
interpolation rule, including its "only when every cell changed" condition --
but the provider arithmetic is a kernel per update on device-resident grids
(csrc/tf_forcing.cu) and the channels component receives R as R1 * t + R0, two
device grids that change only when a provider has updated.  Channel steps
between two provider updates are queued as one tf_chan_step call.

Providers take the reference's "Grid" path: their outputs (P_rain, P_snow, SM,
ET, IN) are grids; every input may be a grid or a scalar.
"""
import ctypes as C

import numpy as np

from . import _lib as L


class _Opd(object):
    """(grid id, scalar) operand of the C ABI."""

    def __init__(self, gid=-1, scalar=0.0):
        self.gid, self.scalar = int(gid), float(scalar)

    def args(self):
        return (self.gid, self.scalar)


class _Provided(object):
    """One provided grid and EMELI's interpolation state for it
    (time_interpolation.py:36-133)."""

    def __init__(self, run, inplace, starts_as_grid=False):
        self.run, self.inplace, self.starts_as_grid = run, inplace, starts_as_grid
        self.obj = run.new_grid()              # the provider's own array
        self.linear = False
        if not inplace:
            self.v1, self.a, self.b = run.new_grid(), run.new_grid(), run.new_grid()
        self.t2 = np.float64(0)

    def live_at_start(self):
        return self.inplace or self.starts_as_grid

    def before_update(self):
        # a provider that rebinds its output leaves the interpolator's v2 untouched,
        # so "v1[:] = v2" (:99) copies the previous value; one that writes in place
        # has already changed v2, v1 becomes the NEW value and nothing is interpolated
        if not self.inplace:
            L.check(L.lib.tf_grid_copy(self.run.h, self.obj, self.v1))

    def after_update(self, t2):
        t1, self.t2 = self.t2, np.float64(t2)
        if self.inplace:
            self.linear = False
            return
        flag = C.c_int(0)
        L.check(L.lib.tf_interp_linear(self.run.h, self.v1, self.obj, float(t1), float(self.t2), self.a, self.b,
                                       C.byref(flag)))
        self.linear = bool(flag.value)

    def value_rate(self):
        """(value operand, rate operand) of a * t + b."""
        if self.linear:
            return _Opd(self.b), _Opd(self.a)
        return _Opd(self.obj), _Opd(-1, 0.0)

    def at(self, time):
        """The operand a user gets at `time` (time_interpolation.py:677)."""
        if not self.linear:
            return _Opd(self.obj)
        tmp = self.run.scratch_grid()
        L.check(L.lib.tf_grid_axpb(self.run.h, self.a, 0.0, self.b, 0.0, float(time), tmp))
        return _Opd(tmp)


class CoupledRun(object):
    """met -> snow (degree-day) -> evap (Priestley-Taylor) -> infil (Green-Ampt)
    -> channels on one strip.  `strip` must be initialised for routing
    (routing.setup_channels); every keyword is a scalar or an [ny, nx] grid.

    met   : P, T_air, T_surf, Qn_SW, Qn_LW, rain_duration, dt_met
    snow  : c0, T0, h0_snow, h0_swe, rho_snow, dt_snow           (None: disabled)
    evap  : alpha, K_soil, soil_x, T_soil_x, dt_evap             (None: disabled)
    infil : Ks, Ki, qs, qi, G, dt_infil [, h_table, elev]        (None: disabled)
    """

    def __init__(self, strip, dt_chan, met, snow=None, evap=None, infil=None, rho_H2O=1000.0):
        self.strip, self.h = strip, strip._h
        self._scratch, self._scratch_next = [], 0
        self.rho_H2O = float(rho_H2O)
        self.time = np.float64(0)
        self.vol = dict(SM=0.0, ET=0.0, IN=0.0, R=0.0)
        self.clock = {}                                        # name -> [time, dt, time_index]
        self.met, self.snow, self.evap, self.infil = dict(met), snow and dict(snow), evap and dict(evap), infil and dict(infil)
        if np.ndim(self.met['P']) == 0 and np.ndim(self.met.get('T_air', 0.0)) == 0:
            raise ValueError('CoupledRun: P or T_air must be a grid (a 0-d P_rain is interpolated '
                             'differently by EMELI; leave all-scalar forcing to the host framework)')
        self.clock['met'] = [np.float64(0), np.float64(self.met['dt']), 0]
        self.P_now = self._operand(self.met['P'])
        self.P0 = self._operand(self.met['P'])
        self.P_rain, self.P_snow = _Provided(self, True), _Provided(self, True)
        for k in ('T_air', 'T_surf', 'Qn_SW', 'Qn_LW'):
            self.met[k] = self._operand(self.met.get(k, 0.0))
        self.SM = self.ET = self.IN = None
        if self.snow:
            self.clock['snow'] = [np.float64(0), np.float64(self.snow['dt']), 0]
            self.SM = _Provided(self, False)
            self.h_snow, self.h_swe = self._state(self.snow['h0_snow']), self._state(self.snow['h0_swe'])
            for k in ('c0', 'T0', 'rho_snow'):
                self.snow[k] = self._operand(self.snow[k])
        if self.evap:
            self.clock['evap'] = [np.float64(0), np.float64(self.evap['dt']), 0]
            self.ET = _Provided(self, False, starts_as_grid=True)       # evap_base: ET starts as a grid of zeros
            for k in ('alpha', 'K_soil', 'soil_x', 'T_soil_x'):
                self.evap[k] = self._operand(self.evap[k])
        if self.infil:
            self.clock['infil'] = [np.float64(0), np.float64(self.infil['dt']), 0]
            self.IN = _Provided(self, False)
            self.I = self._state(1e-6)                         # infil_green_ampt.py:343-351
            for k in ('Ks', 'Ki', 'qs', 'qi', 'G'):
                self.infil[k] = self._operand(self.infil[k])
            self.infil['h_table'] = self._operand(self.infil['h_table']).gid if self.infil.get('h_table') is not None else -1
            self.infil['elev'] = self._operand(self.infil['elev']).gid if self.infil.get('elev') is not None else -1
        self.clock['chan'] = [np.float64(0), np.float64(dt_chan), 0]
        self.order = [n for n in ('met', 'snow', 'evap', 'infil') if n in self.clock] + ['chan']
        self.dt = min(float(c[1]) for c in self.clock.values())
        self._sum0 = self._sum1 = 0.0
        self._pending = 0                                      # channel steps not queued yet
        self.taps, self.bad = [], np.array([-1, -1], dtype=np.int64)
        # time_interpolator.initialize (time_interpolation.py:190-293): one update of every
        # component, users reading their providers' own arrays
        self._init_phase = True
        for name in self.order:
            self._update(name)
        self._flush()
        self._init_phase = False
        self._forcing_dirty = True                             # from now on: interpolated values

    # -- device grids -----------------------------------------------------------------
    def new_grid(self):
        gid = C.c_int(-1)
        L.check(L.lib.tf_grid_new(self.h, C.byref(gid)))
        return gid.value

    def scratch_grid(self):
        if self._scratch_next == len(self._scratch):
            self._scratch.append(self.new_grid())
        self._scratch_next += 1
        return self._scratch[self._scratch_next - 1]

    def _operand(self, v):
        if isinstance(v, _Opd):
            return v
        if np.ndim(v) == 0:
            return _Opd(-1, float(v))
        g = np.ascontiguousarray(v, dtype=np.float64)
        assert g.shape == (self.strip.ny, self.strip.nx), g.shape
        gid = self.new_grid()
        L.check(L.lib.tf_grid_set(self.h, gid, L.ptr(g, C.c_double), 0.0))
        return _Opd(gid)

    def _state(self, v):
        op = self._operand(v)
        if op.gid >= 0:
            return op.gid
        gid = self.new_grid()
        L.check(L.lib.tf_grid_set(self.h, gid, None, op.scalar))
        return gid

    def get(self, gid):
        out = np.empty((self.strip.ny, self.strip.nx), dtype=np.float64)
        L.check(L.lib.tf_grid_get(self.h, int(gid), L.ptr(out, C.c_double)))
        return out

    # -- one component update ------------------------------------------------------------
    def _use(self, prov, time):
        """What a user sees of a provided grid: the provider's own array during
        the initial pass, a * time + b afterwards (emeli.py:1841-1933)."""
        if self._init_phase:
            # initialize_comp_set hands out REFERENCES (emeli.py:1797-1835): a user keeps seeing
            # the array a provider started with.  met writes into that array.  snow and infil
            # start with a 0-d zero and rebind SM / IN at their first update: their users still
            # hold the zero.  evap starts with a GRID of zeros, which the interpolator aliases and
            # overwrites in place with the new values (time_interpolation.py:73, :107): live too.
            return _Opd(prov.obj) if prov.live_at_start() else _Opd(-1, 0.0)
        return prov.at(time)

    def _update(self, name):
        clk = self.clock[name]
        t = clk[0]
        self._scratch_next = 0
        if name == 'chan':
            if self._forcing_dirty:
                self._flush()
                self._hand_over_forcing()
            self._pending += 1
            self.vol['R'] += self._sum1 * float(t) + self._sum0              # update_R_integral
        else:
            self._flush()                                       # queued channel steps read the old grids
            getattr(self, '_update_' + name)(t, float(clk[1]), clk[2])
            self._forcing_dirty = True
        clk[2] += 1
        clk[0] = clk[0] + clk[1]                                # BMI_base.update_time :1067-1114
        for prov in {'met': (self.P_rain, self.P_snow), 'snow': (self.SM,), 'evap': (self.ET,),
                     'infil': (self.IN,)}.get(name, ()):
            prov.after_update(clk[0])

    _forcing_dirty = True

    def _update_met(self, t, dt, index):
        m = self.met
        L.check(L.lib.tf_prov_met_rain(self.h, *self.P_now.args(), *m['T_air'].args(), self.P_rain.obj, self.P_snow.obj))
        if index > 0:                                           # read_input_files: the next entry of the sequence
            on = float(t + dt) < float(m.get('rain_duration', np.inf))
            self.P_now = self.P0 if on else _Opd(-1, 0.0)

    def _update_snow(self, t, dt, index):
        s, m = self.snow, self.met
        self.SM.before_update()
        vol = C.c_double(0.0)
        L.check(L.lib.tf_prov_snow_degree_day(self.h, *m['T_air'].args(), *self._use(self.P_snow, t).args(),
                                              *s['c0'].args(), *s['T0'].args(), self.rho_H2O, *s['rho_snow'].args(),
                                              dt, self.h_snow, self.h_swe, self.SM.obj, C.byref(vol)))
        self.vol['SM'] += vol.value

    def _update_evap(self, t, dt, index):
        e, m = self.evap, self.met
        self.ET.before_update()
        vol = C.c_double(0.0)
        L.check(L.lib.tf_prov_evap_priestley_taylor(self.h, *m['Qn_SW'].args(), *m['Qn_LW'].args(), *m['T_air'].args(),
                                                    *m['T_surf'].args(), *e['alpha'].args(), *e['K_soil'].args(),
                                                    *e['soil_x'].args(), *e['T_soil_x'].args(), dt, self.ET.obj,
                                                    C.byref(vol)))
        self.vol['ET'] += vol.value

    def _update_infil(self, t, dt, index):
        f = self.infil
        sm = self._use(self.SM, t) if self.SM is not None else _Opd(-1, 0.0)
        self.IN.before_update()
        vol = C.c_double(0.0)
        L.check(L.lib.tf_prov_infil_green_ampt(self.h, *self._use(self.P_rain, t).args(), *sm.args(), *f['Ks'].args(),
                                               *f['Ki'].args(), *f['qs'].args(), *f['qi'].args(), *f['G'].args(),
                                               f['h_table'], f['elev'], dt, self.I, self.IN.obj, C.byref(vol)))
        self.vol['IN'] += vol.value

    # -- channels ----------------------------------------------------------------------------
    def _hand_over_forcing(self):
        """update_R's inputs as value + rate * t (channels_base.py:1103-1147)."""
        zero = (_Opd(-1, 0.0), _Opd(-1, 0.0))
        terms = {L.R_P_RAIN: self.P_rain, L.R_SM: self.SM, L.R_GW: None, L.R_MR: None, L.R_ET: self.ET,
                 L.R_IN: self.IN}
        for which, prov in terms.items():
            if prov is None:
                v, r = zero
            elif self._init_phase:
                v, r = (_Opd(prov.obj) if prov.live_at_start() else _Opd(-1, 0.0)), _Opd(-1, 0.0)
            else:
                v, r = prov.value_rate()
            L.check(L.lib.tf_chan_set_forcing_dev(self.h, int(which), *v.args(), *r.args()))
        s0, s1 = C.c_double(0.0), C.c_double(0.0)
        L.check(L.lib.tf_chan_assemble_R_dev(self.h, C.byref(s0), C.byref(s1)))
        self._sum0, self._sum1 = s0.value, s1.value
        self._forcing_dirty = False

    def _flush(self):
        """Queue the channel steps counted so far as one call."""
        if self._pending:
            taps, _, bad = self.strip.step(self._pending)
            self.taps.append(taps)
            for k in (0, 1):
                if bad[k] != -1 and self.bad[k] == -1:
                    self.bad[k] = bad[k]
            self._pending = 0

    # -- the framework loop --------------------------------------------------------------------
    def sweep(self):
        """One pass of EMELI's while-loop body (emeli.py:1359-1418)."""
        for name in self.order:
            if self.time > self.clock[name][0]:
                self._update(name)
        self.time = self.time + np.float64(self.dt)

    def run_until(self, seconds):
        while float(self.clock['chan'][0]) < seconds:
            self.sweep()
        self._flush()

    def hydrograph(self):
        """Outlet taps [n_steps, n_outlets, 4] of every channel step so far."""
        self._flush()
        return np.concatenate(self.taps, axis=0) if self.taps else np.zeros((0, self.strip.n_outlets, 4))
