"""Host-side initialisation of the device routing state.

`setup_channels` performs, in NumPy and once per run, what the reference does
in channels_component.initialize_computed_vars (channels_base.py:797-984) and
hands the results to the CUDA library.  It is shared by the BMI components
(topoflow_old_b200/components/channels_base.py), bench.py and the tests.
"""
import numpy as np

from . import _lib as L
from .engine import Strip

G = np.float64(9.81)


def remove_bad_slopes(slope, report=None, reduce_min=None):
    """channels_base.py:2675-2754: zero, negative and non-finite slopes take
    the smallest positive slope of the grid.  Returns the float64 grid."""
    wz = (slope == 0.0)
    wn = (slope < 0.0)
    wi = np.logical_not(np.isfinite(slope))
    good = np.invert(np.logical_or(slope <= 0.0, wi))
    nz, nn, ni = int(wz.sum()), int(wn.sum()), int(wi.sum())
    if reduce_min is None and ((nz + nn + ni) == 0 or not good.any()):
        return slope
    s_min = slope[good].min() if good.any() else np.inf
    if reduce_min is not None:
        # several strips: the smallest positive slope of the WHOLE grid
        s_min = reduce_min(float(s_min))
        if (nz + nn + ni) == 0 or not np.isfinite(s_min):
            return slope
    if report is not None:
        report.update(nz=nz, nn=nn, ni=ni, S_min=s_min,
                      S_max=slope[good].max() if good.any() else s_min)
    slope[wz] = s_min
    slope[wn] = s_min
    slope[wi] = s_min
    return np.float64(slope)


def channel_ds(sinu, ds):
    """d8.ds = sinu * d8.ds (channels_base.py:882).  Under the NumPy 1.x
    value-based casting the reference ran on, a 0-d sinuosity leaves the
    float32 flow lengths float32."""
    if np.ndim(sinu) == 0:
        return np.float32(sinu) * ds
    return sinu * ds


def setup_channels(strip, wave, law, dt, da, slope, width, angle, rough,
                   sinu=1.0, d0=0.0, outlets=(), allow_packed=True,
                   ds=None, report=None, reduce_min=None):
    """Upload parameters and the initial state; returns a dict of the host
    arrays the BMI component keeps (S_bed, ds, initial vol ...).

    slope, width, angle [rad], rough (nval or z0val), sinu, d0: 0-d or
    [ny, nx] float64, as they stand after read_input_files()
    (channels_base.py:2142-2200).  The D8 grid must already be on the strip.
    reduce_min: with several strips, a callable returning the minimum of a
    float over all strips (remove_bad_slopes needs the grid-wide value).
    """
    assert isinstance(strip, Strip)
    ny, nx = strip.ny, strip.nx
    if ds is None:
        ds = strip.d8_get(L.D8_DS)
    ds = channel_ds(sinu, ds)                                   # :882
    slope = slope / sinu                                        # :887
    if np.ndim(slope) == 0:
        slope = np.zeros((ny, nx), dtype='float64') + slope
    slope = np.float64(slope)
    if wave == L.KINEMATIC:
        slope = remove_bad_slopes(slope, report, reduce_min)    # :949-950
    d = np.zeros((ny, nx), dtype='float64')
    dry = np.ndim(d0) == 0 and float(d0) == 0.0
    if dry:                                                     # the usual start: nothing to compute
        A_wet = np.zeros((ny, nx), dtype='float64')
        vol = np.zeros((ny, nx), dtype='float64')
    else:
        d += d0                                                 # :903
        L2 = d * np.tan(angle)                                  # :957-960
        A_wet = d * (width + L2)
        vol = A_wet * ds
    strip.set_param(L.P_WIDTH, width, is_grid=np.ndim(width) > 0)
    # tan / cos with NumPy (the reference's libm), but only of the distinct angles
    if not strip.set_angle(angle, is_grid=np.ndim(angle) > 0):
        strip.set_param(L.P_TAN_ANGLE, np.tan(angle), is_grid=np.ndim(angle) > 0)
        strip.set_param(L.P_COS_ANGLE, np.cos(angle), is_grid=np.ndim(angle) > 0)
    strip.set_param(L.P_SLOPE, slope, is_grid=True)
    strip.set_param(L.P_ROUGH, rough, is_grid=np.ndim(rough) > 0)
    strip.set_param(L.P_DS, np.float64(ds), is_grid=True)
    strip.chan_init(wave, law, dt, da, outlets=outlets, allow_packed=allow_packed)
    if not dry:                                                 # (tf_chan_init leaves vol = d = 0)
        strip.set_state(L.F_VOL, vol)
        strip.set_state(L.F_D, d)
    # u = 0, so the first Q = u * A_wet is 0 everywhere (:900, :1188)
    return dict(S_bed=slope, ds=ds, vol0=vol, d0=d, A_wet0=A_wet)
