"""Seeded test cases shared by the GPU parity tests: the same inputs are fed
to the oracle (oracle/) and to the CUDA path through the C ABI."""
import numpy as np

from oracle import channels as och
from oracle import d8 as od8
from topoflow_old_b200.utils import synth

DEG = np.pi / 180.0


def make_case(nx, ny, seed=7, cell=30.0, angle_mode='classes', exact_f32=True,
              scalar_params=False, sinu=1.0, d0=0.0, rain_mmph=60.0, rain_grid=False,
              roughness=8.0, law=0):
    rng = np.random.default_rng(seed)
    dem = synth.fractal_dem(nx, ny, seed=seed, cell=cell, roughness=roughness)
    # a flat patch and a nodata hole so link_flats and the code-0 paths run
    if nx > 24 and ny > 24:
        dem[ny // 3:ny // 3 + 4, nx // 3:nx // 3 + 5] = dem[ny // 3, nx // 3]
        dem[2 * ny // 3, 2 * nx // 3] = np.float32(-9999.0)
    dx, dy, dd = od8.pixel_sizes(ny, cell, cell)
    b = od8.build(dem, cell, cell)
    slope = synth.d8_slope(dem, b['d8_grid'], b['ds'])
    c = dict(nx=nx, ny=ny, dem=dem, dx=dx, dy=dy, dd=dd, cell=cell, d8=b,
             da=np.float64(cell) * np.float64(cell), sinu=sinu, d0=d0, law=law)
    if scalar_params:
        c['width'] = och.scalar(3.0)
        c['angle'] = och.scalar(0.5 if angle_mode != 'zero' else 0.0)
        c['rough'] = och.scalar(0.05 if law == 0 else 0.01)
    else:
        w = synth.uniform_field(nx, ny, 1.0, 6.0, seed=seed + 1)
        w[rng.random((ny, nx)) < 0.02] = 0.0               # "width == 0 -> dw"
        if angle_mode == 'zero':
            a = np.zeros((ny, nx), dtype=np.float32)
        elif angle_mode == 'classes':
            a = synth.class_field(nx, ny, [0.0, 20.0, 30.0, 45.0, 60.0], seed=seed + 2, block=5)
        else:                                              # every cell its own angle
            a = synth.uniform_field(nx, ny, 5.0, 70.0, seed=seed + 2)
        r = (synth.uniform_field(nx, ny, 0.03, 0.15, seed=seed + 3) if law == 0 else
             synth.uniform_field(nx, ny, 0.001, 0.05, seed=seed + 3))
        if not exact_f32:                                  # float64-only values
            w = np.float64(w) * (1.0 + 1e-9)
            r = np.float64(r) * (1.0 + 1e-9)
        inp = och.prepare_inputs(b, slope, np.float64(w), a, nval=np.float64(r))
        c['width'], c['angle'], c['rough'] = inp['width'], inp['angle'], inp['nval']
    c['slope'] = np.float64(slope)
    if not exact_f32:
        c['slope'] = c['slope'] * (1.0 + 1e-9)
    # outlet: the interior cell with the largest contributing area + two more
    A = od8.area_grid_fast(b['d8_grid'], c['da'])
    A[0, :] = A[-1, :] = 0
    A[:, 0] = A[:, -1] = 0
    o = np.unravel_index(np.argmax(A), A.shape)
    c['outlets'] = [(int(o[0]), int(o[1])), (ny // 2, nx // 2), (ny // 4, (3 * nx) // 4)]
    P = rain_mmph / 3.6e6
    if rain_grid:
        c['P_rain'] = P * (0.5 + synth.uniform_field(nx, ny, 0.0, 1.0, seed=seed + 4, dtype=np.float64))
    else:
        c['P_rain'] = och.scalar(P)
    return c


def oracle_for(c, wave, dt):
    cp = lambda v: v.copy() if np.ndim(v) else och.scalar(float(v))
    law = c['law']
    ch = och.ChannelsOracle(c['d8'], wave, dt=dt, da=c['da'], slope=c['slope'].copy(),
                            width=cp(c['width']), angle=cp(c['angle']),
                            nval=cp(c['rough']) if law == 0 else None,
                            z0val=cp(c['rough']) if law == 1 else None,
                            sinu=c['sinu'], d0=c['d0'], manning=(law == 0),
                            law_of_wall=(law == 1), outlet_ID=c['outlets'][0])
    ch.P_rain = c['P_rain']
    return ch


def rel_err(got, want, floor=1e-300):
    """max |got - want| / max(|want|, floor), elementwise; NaNs must coincide."""
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    m = ~np.isnan(want)
    if not m.any():
        return 0.0
    den = np.maximum(np.abs(want[m]), floor)
    return float(np.max(np.abs(got[m] - want[m]) / den))
