"""Seeded synthetic inputs for the routing path (bench.py and the parity tests).

Everything is a pure function of (seed, global row, global column), so any
row strip of a grid can be generated on its own and strips of different ranks
fit together.  The DEM is a tilted plane (draining to the south-east corner)
plus fractal value noise whose total gradient stays below the tilt, so the
surface has no pits by construction and every cell drains to the border --
the generator needs no pit filling, which the reference would do with a
serial priority flood (utils/fill_pits.py).
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _hash01(ix, iy, salt):
    """Uniform [0, 1) float64 from integer lattice coordinates (splitmix64)."""
    with np.errstate(over='ignore'):
        z = (ix.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)
             + iy.astype(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F)
             + np.uint64(salt) * np.uint64(0x165667B19E3779F9))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def _value_noise(rows, cols, spacing, salt):
    """Smooth noise in [0, 1], lattice `spacing` cells, on rows x cols."""
    fy = rows.astype(np.float64) / spacing
    fx = cols.astype(np.float64) / spacing
    iy0 = np.floor(fy).astype(np.int64)
    ix0 = np.floor(fx).astype(np.int64)
    ty = (fy - iy0)[:, None]
    tx = (fx - ix0)[None, :]
    ty = ty * ty * (3.0 - 2.0 * ty)
    tx = tx * tx * (3.0 - 2.0 * tx)
    # lattice values on the (few) distinct lattice rows / columns
    uy, inv_y = np.unique(np.concatenate([iy0, iy0 + 1]), return_inverse=True)
    ux, inv_x = np.unique(np.concatenate([ix0, ix0 + 1]), return_inverse=True)
    lat = _hash01(ux[None, :], uy[:, None], salt)
    ny, nx = rows.size, cols.size
    y0, y1 = inv_y[:ny], inv_y[ny:]
    x0, x1 = inv_x[:nx], inv_x[nx:]
    top = lat[y0][:, x0] * (1.0 - tx) + lat[y0][:, x1] * tx
    bot = lat[y1][:, x0] * (1.0 - tx) + lat[y1][:, x1] * tx
    return top * (1.0 - ty) + bot * ty


def fractal_dem(nx, ny, row0=0, nrows=None, seed=20170226, cell=30.0, tilt=0.01,
                hurst=0.8, roughness=0.85):
    """float32 [nrows, nx] rows row0 .. row0+nrows of an ny x nx DEM [m].

    tilt      : slope of the regional plane along each axis [m/m]
    roughness : fraction (< 1) of the tilt that the noise gradient may reach
    """
    nrows = ny - row0 if nrows is None else nrows
    rows = np.arange(row0, row0 + nrows)
    cols = np.arange(nx)
    tc = tilt * cell                                  # drop per cell along one axis
    z = tc * ((ny - 1 - rows)[:, None] + (nx - 1 - cols)[None, :]).astype(np.float64)
    spacings = [s for s in (4, 8, 16, 32, 64, 128, 256, 512, 1024) if s <= 2 * max(nx, ny)]
    lmax = float(spacings[-1])
    # smoothstep value noise: each partial derivative is <= 1.5 * amp / spacing,
    # so a step to the south-east neighbour changes the noise by at most
    # 2 * amp * grad while the plane drops 2 * tc: amp * grad < tc keeps that
    # neighbour strictly lower, hence no pits
    weights = [(s / lmax) ** hurst for s in spacings]
    grad = sum(1.5 * w / s for w, s in zip(weights, spacings))
    amp = roughness * tc / grad
    for k, (s, w) in enumerate(zip(spacings, weights)):
        z += amp * w * _value_noise(rows, cols, float(s), seed + 101 * k)
    return z.astype(np.float32)


def uniform_field(nx, ny, lo, hi, row0=0, nrows=None, seed=1, dtype=np.float32):
    """Per-cell uniform values in [lo, hi) (e.g. Manning n), as `dtype`."""
    nrows = ny - row0 if nrows is None else nrows
    rows = np.arange(row0, row0 + nrows, dtype=np.int64)
    cols = np.arange(nx, dtype=np.int64)
    u = _hash01(cols[None, :], rows[:, None], seed)
    return (lo + (hi - lo) * u).astype(dtype)


def class_field(nx, ny, values, row0=0, nrows=None, seed=2, block=16, dtype=np.float32):
    """Piecewise-constant field: each block x block patch takes one of `values`
    (e.g. bank angles by reach class)."""
    nrows = ny - row0 if nrows is None else nrows
    rows = np.arange(row0, row0 + nrows, dtype=np.int64) // block
    cols = np.arange(nx, dtype=np.int64) // block
    u = _hash01(cols[None, :], rows[:, None], seed)
    idx = np.minimum((u * len(values)).astype(np.int64), len(values) - 1)
    return np.asarray(values, dtype=dtype)[idx]


def d8_slope(dem, codes, ds, floor=1e-5, north=None, south=None):
    """float32 bed slope = drop to the D8 parent / flow length, clipped to
    >= floor; 0 where the cell has no flow direction (the routing init then
    applies the reference's remove_bad_slopes).  north / south: the halo DEM
    rows of a strip."""
    ny, nx = dem.shape
    z = np.empty((ny + 2, nx + 2), dtype=np.float64)
    z[1:-1, 1:-1] = dem
    z[0, 1:-1] = dem[0] if north is None else north
    z[-1, 1:-1] = dem[-1] if south is None else south
    z[:, 0] = z[:, 1]
    z[:, -1] = z[:, -2]
    out = np.zeros((ny, nx), dtype=np.float64)
    nbr = ((-1, 1), (0, 1), (1, 1), (1, 0), (1, -1), (0, -1), (-1, -1), (-1, 0))
    for k, (dr, dc) in enumerate(nbr):
        m = codes == (1 << k)
        if m.any():
            zn = z[1 + dr:1 + dr + ny, 1 + dc:1 + dc + nx]
            out[m] = (dem[m].astype(np.float64) - zn[m]) / ds[m]
    flowing = codes != 0
    out[flowing] = np.maximum(out[flowing], floor)
    return out.astype(np.float32)
