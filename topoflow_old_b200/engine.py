"""NumPy-facing wrapper of one device strip (a tf_handle).

This is the only module that talks to the C ABI; the BMI components in
topoflow_old_b200/components build on it.  All arrays passed in and out are
host NumPy arrays; the library copies them.
"""
import ctypes as C

import numpy as np

from . import _lib as L

_D8_DTYPES = {L.D8_CODE: np.uint8, L.D8_PARENT_ID: np.int32, L.D8_DS: np.float32,
              L.D8_DW: np.float32, L.D8_AREA: np.float32, L.D8_INMASK: np.uint8,
              L.D8_RAW: np.int16}


def device_count():
    return L.lib.tf_device_count()


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class Strip(object):
    """ny x nx cells on one GPU; the whole grid unless set_strip() places it
    inside a taller one."""

    def __init__(self, nx, ny, device=0):
        self.nx, self.ny = int(nx), int(ny)
        self._h = L.handle_t()
        L.check(L.lib.tf_create(self.nx, self.ny, int(device), 0, C.byref(self._h)))
        self.n_outlets = 0
        self.n_canals = 0
        self._pinned = []

    # -- lifetime ---------------------------------------------------------
    def close(self):
        if getattr(self, '_h', None) is not None and self._h.value:
            L.lib.tf_destroy(self._h)          # waits for a frame copy in flight
            self._h = L.handle_t()
            for p in self._pinned:
                L.lib.tf_host_free(p)
            self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def sync(self):
        L.check(L.lib.tf_sync(self._h))

    def set_strip(self, row_offset, global_ny, has_north, has_south):
        L.check(L.lib.tf_set_strip(self._h, int(row_offset), int(global_ny),
                                   int(bool(has_north)), int(bool(has_south))))

    # -- D8 -----------------------------------------------------------------
    def _rows(self, dx, dy, dd):
        out = []
        for a in (dx, dy, dd):
            a = np.ascontiguousarray(np.broadcast_to(np.float32(a), (self.ny,)), dtype=np.float32)
            out.append(a)
        return out

    def d8_build(self, dem, dx, dy, dd, nodata=-9999.0, link_flats=True,
                 lr_periodic=False, tb_periodic=False):
        dem = np.ascontiguousarray(dem, dtype=np.float32)
        assert dem.shape == (self.ny, self.nx), dem.shape
        dx, dy, dd = self._rows(dx, dy, dd)
        stats = np.zeros(4, dtype=np.int64)
        L.check(L.lib.tf_d8_build(self._h, L.ptr(dem, C.c_float), L.ptr(dx, C.c_float),
                                  L.ptr(dy, C.c_float), L.ptr(dd, C.c_float),
                                  C.c_float(nodata), int(bool(link_flats)),
                                  int(bool(lr_periodic)), int(bool(tb_periodic)),
                                  L.ptr(stats, C.c_int64)))
        return dict(n_flats=int(stats[0]), n_pits=int(stats[1]), n_bad=int(stats[2]),
                    n_reps=int(stats[3]))

    def d8_set_row0_sizes(self, dx0, dy0, dd0):
        """dx, dy, dd of row 0 of the global grid (d8_global.py:209-211) for a
        strip that does not hold that row."""
        L.check(L.lib.tf_d8_set_row0_sizes(self._h, C.c_float(dx0), C.c_float(dy0), C.c_float(dd0)))

    def d8_set_codes(self, codes, dx, dy, dd, lr_periodic=False, tb_periodic=False):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        assert codes.shape == (self.ny, self.nx)
        dx, dy, dd = self._rows(dx, dy, dd)
        L.check(L.lib.tf_d8_set_codes(self._h, L.ptr(codes, C.c_uint8), L.ptr(dx, C.c_float),
                                      L.ptr(dy, C.c_float), L.ptr(dd, C.c_float),
                                      int(bool(lr_periodic)), int(bool(tb_periodic))))

    def d8_set_halo_dem(self, north=None, south=None):
        n = None if north is None else np.ascontiguousarray(north, dtype=np.float32)
        s = None if south is None else np.ascontiguousarray(south, dtype=np.float32)
        L.check(L.lib.tf_d8_set_halo_dem(self._h, L.ptr(n, C.c_float), L.ptr(s, C.c_float)))

    def d8_area(self, da, units_km=True):
        da_rows = np.ascontiguousarray(np.broadcast_to(np.float64(da), (self.ny,)), dtype=np.float64)
        n_bad = C.c_int64(0)
        L.check(L.lib.tf_d8_area(self._h, L.ptr(da_rows, C.c_double), int(bool(units_km)),
                                 C.byref(n_bad)))
        self.area_rounds = int(L.lib.tf_d8_area_rounds(self._h))
        return int(n_bad.value)

    def d8_get(self, which):
        out = np.empty((self.ny, self.nx), dtype=_D8_DTYPES[which])
        L.check(L.lib.tf_d8_get(self._h, int(which), out.ctypes.data_as(C.c_void_p)))
        return out

    def fill_pits(self, dem, nodata=-9999.0):
        """utils/fill_pits.py fill_pits() on the device: (filled DEM float32,
        number of raised cells, passes over the grid)."""
        dem = np.ascontiguousarray(dem, dtype=np.float32)
        assert dem.shape == (self.ny, self.nx), dem.shape
        out = np.empty_like(dem)
        n_raised, n_sweeps = C.c_int64(0), C.c_int64(0)
        L.check(L.lib.tf_fill_pits(self._h, L.ptr(dem, C.c_float), C.c_float(nodata), L.ptr(out, C.c_float),
                                   C.byref(n_raised), C.byref(n_sweeps)))
        return out, int(n_raised.value), int(n_sweeps.value)

    def satzone_dzw(self, Q_gw, Rg, da, dt):
        """Water-table change from recharge and lateral D8 flow
        (satzone_base.py:930-1008)."""
        Q = _f64(Q_gw)
        assert Q.shape == (self.ny, self.nx)
        da_rows = np.ascontiguousarray(np.broadcast_to(np.float64(da), (self.ny,)), dtype=np.float64)
        out = np.empty_like(Q)
        if np.ndim(Rg) == 0:
            L.check(L.lib.tf_satzone_dzw(self._h, L.ptr(Q, C.c_double), None, float(Rg), L.ptr(da_rows, C.c_double),
                                         float(dt), L.ptr(out, C.c_double)))
        else:
            R = _f64(Rg)
            L.check(L.lib.tf_satzone_dzw(self._h, L.ptr(Q, C.c_double), L.ptr(R, C.c_double), 0.0,
                                         L.ptr(da_rows, C.c_double), float(dt), L.ptr(out, C.c_double)))
        return out

    # -- channels -------------------------------------------------------------
    def set_param(self, which, value, is_grid=None):
        """value: 0-d / Python scalar or [ny, nx] array."""
        if np.ndim(value) == 0:
            L.check(L.lib.tf_chan_set_param(self._h, int(which), None, float(value),
                                            int(bool(is_grid))))
        else:
            g = _f64(value)
            assert g.shape == (self.ny, self.nx), g.shape
            L.check(L.lib.tf_chan_set_param(self._h, int(which), L.ptr(g, C.c_double), 0.0,
                                            1 if is_grid is None else int(bool(is_grid))))

    def set_angle(self, angle, is_grid=None):
        """Hand over the bank angle [rad] (0-d or [ny, nx]); tan and cos are
        evaluated here, with NumPy, for the distinct values only.  Returns False
        when the grid holds more than 255 distinct angles (then use set_param
        with TF_P_TAN_ANGLE / TF_P_COS_ANGLE grids)."""
        distinct = np.zeros(255, dtype=np.float64)
        n = C.c_int(-1)
        if np.ndim(angle) == 0:
            L.check(L.lib.tf_chan_set_angle(self._h, None, float(angle), int(bool(is_grid)),
                                            L.ptr(distinct, C.c_double), C.byref(n)))
        else:
            g = _f64(angle)
            assert g.shape == (self.ny, self.nx), g.shape
            L.check(L.lib.tf_chan_set_angle(self._h, L.ptr(g, C.c_double), 0.0,
                                            1 if is_grid is None else int(bool(is_grid)),
                                            L.ptr(distinct, C.c_double), C.byref(n)))
        if n.value < 0:
            return False
        a = np.ascontiguousarray(distinct[:n.value])
        t, c = np.ascontiguousarray(np.tan(a)), np.ascontiguousarray(np.cos(a))
        L.check(L.lib.tf_chan_set_angle_table(self._h, L.ptr(t, C.c_double), L.ptr(c, C.c_double), n.value))
        return True

    def chan_init(self, wave, law, dt, da, outlets=(), allow_packed=True):
        da_rows = np.ascontiguousarray(np.broadcast_to(np.float64(da), (self.ny,)), dtype=np.float64)
        rc = np.ascontiguousarray(np.array(outlets, dtype=np.int32).reshape(-1, 2))
        self.n_outlets = rc.shape[0]
        L.check(L.lib.tf_chan_init(self._h, int(wave), int(law), float(dt),
                                   L.ptr(da_rows, C.c_double), L.ptr(rc, C.c_int32),
                                   self.n_outlets, int(bool(allow_packed))))

    def set_rho(self, rho):
        L.check(L.lib.tf_chan_set_rho(self._h, float(rho)))

    def set_state(self, which, grid):
        g = _f64(grid)
        assert g.shape == (self.ny, self.nx)
        L.check(L.lib.tf_chan_set_state(self._h, int(which), L.ptr(g, C.c_double)))

    def set_R(self, R, want_sum=True):
        """Returns sum over the strip of (R*da)*dt (the vol_R increment)."""
        s = C.c_double(0.0)
        sp = C.byref(s) if want_sum else None
        if np.ndim(R) == 0:
            L.check(L.lib.tf_chan_set_R(self._h, None, float(R), sp))
        else:
            g = _f64(R)
            assert g.shape == (self.ny, self.nx)
            L.check(L.lib.tf_chan_set_R(self._h, L.ptr(g, C.c_double), 0.0, sp))
        return s.value

    def set_forcing(self, which, value):
        """One of the six fluxes of update_R (L.R_P_RAIN ...): 0-d or [ny, nx]."""
        if np.ndim(value) == 0 or np.size(value) == 1:
            L.check(L.lib.tf_chan_set_forcing(self._h, int(which), None, float(np.asarray(value).reshape(-1)[0])))
        else:
            g = _f64(value)
            assert g.shape == (self.ny, self.nx)
            L.check(L.lib.tf_chan_set_forcing(self._h, int(which), L.ptr(g, C.c_double), 0.0))

    def assemble_R(self, want_sum=True):
        """R = (P_rain + SM + GW + MR) - (ET + IN) on the device; returns the
        vol_R increment sum((R*da)*dt) when asked."""
        out = C.c_double(0.0)
        L.check(L.lib.tf_chan_assemble_R(self._h, C.byref(out) if want_sum else None))
        return out.value if want_sum else None

    def set_points(self, kind, rows, cols, values):
        rows = np.ascontiguousarray(np.atleast_1d(rows), dtype=np.int32)
        cols = np.ascontiguousarray(np.atleast_1d(cols), dtype=np.int32)
        vals = np.ascontiguousarray(np.broadcast_to(np.asarray(values, dtype=np.float64), rows.shape))
        if kind == L.PT_CANALS_IN:
            self.n_canals = rows.size
        L.check(L.lib.tf_chan_set_points(self._h, int(kind), L.ptr(rows, C.c_int32),
                                         L.ptr(cols, C.c_int32), L.ptr(vals, C.c_double),
                                         rows.size))

    def step(self, n_steps=1):
        """Returns (taps [n_steps, n_outlets, 4], canals_in [n_steps, n_canals],
        bad [2]) -- bad = first bad-depth / bad-velocity cell index or -1."""
        taps = np.zeros((n_steps, self.n_outlets, 4), dtype=np.float64)
        cin = np.zeros((n_steps, self.n_canals), dtype=np.float64)
        bad = np.zeros(2, dtype=np.int64)
        L.check(L.lib.tf_chan_step(self._h, int(n_steps),
                                   L.ptr(taps, C.c_double) if self.n_outlets else None,
                                   L.ptr(cin, C.c_double) if self.n_canals else None,
                                   L.ptr(bad, C.c_int64)))
        return taps, cin, bad

    def step_async(self, n_steps=1):
        L.check(L.lib.tf_chan_step_async(self._h, int(n_steps)))

    def get(self, which):
        out = np.empty((self.ny, self.nx), dtype=np.float64)
        L.check(L.lib.tf_chan_get(self._h, int(which), L.ptr(out, C.c_double)))
        return out

    def frame(self, which, swap_endian=False, out=None, wait=True):
        """One field as the float32 [ny, nx] frame the reference appends to an
        .rts stack (rts_files.py:313-396), rounded and byte-swapped on the
        device.  With wait=False the copy is still in flight on return; call
        frame_wait() before reading `out` (give a pinned_frame() for a truly
        asynchronous copy)."""
        if out is None:
            out = np.empty((self.ny, self.nx), dtype=np.float32)
        assert out.dtype == np.float32 and out.size == self.ny * self.nx and out.flags.c_contiguous
        L.check(L.lib.tf_chan_frame(self._h, int(which), int(bool(swap_endian)),
                                    L.ptr(out, C.c_float), 0 if wait else 1))
        return out

    def frame_wait(self):
        L.check(L.lib.tf_chan_frame_wait(self._h))

    def pinned_array(self, shape, dtype):
        """A page-locked host array (freed with the strip): frames, forcing grids."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape))
        p = C.c_void_p()
        L.check(L.lib.tf_host_alloc(n * dtype.itemsize, C.byref(p)))
        self._pinned.append(p)
        buf = (C.c_char * (n * dtype.itemsize)).from_address(p.value)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    def pinned_frame(self):
        """A page-locked float32 [ny, nx] array for frame()."""
        return self.pinned_array((self.ny, self.nx), np.float32)

    def minmax(self):
        out = np.zeros(6, dtype=np.float64)
        L.check(L.lib.tf_chan_minmax(self._h, L.ptr(out, C.c_double)))
        return out

    def timing(self):
        """(elapsed_ms, steps, launches, layout bytes per cell-update) since
        the previous call."""
        out = np.zeros(4, dtype=np.float64)
        L.check(L.lib.tf_chan_timing(self._h, L.ptr(out, C.c_double)))
        return out

    # -- strips over several GPUs --------------------------------------------
    @staticmethod
    def comm_unique_id():
        buf = np.zeros(128, dtype=np.uint8)
        L.check(L.lib.tf_comm_unique_id(L.ptr(buf, C.c_uint8)))
        return buf

    def comm_init(self, rank, nranks, unique_id):
        buf = np.ascontiguousarray(unique_id, dtype=np.uint8)
        L.check(L.lib.tf_comm_init(self._h, int(rank), int(nranks), L.ptr(buf, C.c_uint8)))
