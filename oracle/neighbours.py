"""TEST INFRASTRUCTURE (oracle): restatements of two users of the D8 structures
next to the routing path -- utils/fill_pits.py and the lateral-flow part of
components/satzone_base.py update_water_table.  Parity UNPINNED: the reference
ships no output of either (FILL_PITS_IN_Z0 is off in every shipped cfg; Test10's
satzone run cannot be replayed without the Darcy-layer component), so these are
restatements of the cited lines, used on small grids only."""
import heapq

import numpy as np


def start_pixels(DEM, nodata):
    """get_start_pixels(), utils/fill_pits.py:178-330 (EDGES_ONLY False).  DEM is
    changed in place: non-finite cells become nodata (:239-242)."""
    ny, nx = DEM.shape
    DEM[~np.isfinite(DEM)] = nodata
    ids = []
    ids += list(np.where(DEM[0, :] > nodata)[0])
    ids += list(np.where(DEM[1:ny - 1, 0] > nodata)[0] * nx)           # (:258: rows counted from 1, as there)
    ids += list(np.where(DEM[1:ny - 1, nx - 1] > nodata)[0] * nx + (nx - 1))
    ids += list(np.where(DEM[ny - 1, :] > nodata)[0] + nx * (ny - 1))
    zc = DEM[1:ny - 1, 1:nx - 1]
    zs = np.full(zc.shape, np.inf, dtype=DEM.dtype)
    for dy, dx in ((1, -1), (0, -1), (-1, -1), (-1, 0), (-1, 1), (0, 1), (1, 1), (1, 0)):
        zs = np.minimum(zs, np.roll(np.roll(zc, dy, axis=0), dx, axis=1))
    w = np.where(np.logical_and(zc.flat > nodata, zs.flat <= nodata))[0]
    # DEVIATION, on purpose: the reference turns these positions into cell IDs with "w + nx + 1"
    # (:322), i.e. it uses the flat index of the INTERIOR array (nx - 2 columns) as if it were one
    # of the full grid, which puts every such start cell of interior row r two r columns too far
    # west.  The restatement (and the CUDA path) use the cells the comment there describes.
    r, c = np.divmod(w, nx - 2)
    ids += list((r + 1) * nx + (c + 1))
    return np.array(ids, dtype=np.int64)


def fill_pits(DEM, nodata=np.float32(-9999)):
    """fill_pits(), utils/fill_pits.py:331-582 with heapq (USE_MY_HEAP False).
    Returns (filled DEM, n_raised)."""
    DEM = np.array(DEM, dtype=np.float32)
    ny, nx = DEM.shape
    n = nx * ny
    IDs = start_pixels(DEM, nodata)
    OPEN, CLOSED, ON_HEAP = 0, 1, 2
    C = np.zeros(n, dtype=np.uint8)
    flat = DEM.reshape(-1)
    C[np.logical_or(flat <= nodata, ~np.isfinite(flat))] = CLOSED
    heap = [(float(flat[i]), int(i)) for i in IDs]
    heapq.heapify(heap)
    C[IDs] = ON_HEAP
    incs = (-nx - 1, -nx, -nx + 1, -1, 1, nx - 1, nx, nx + 1)
    n_raised = 0
    while heap:
        zmin, i = heapq.heappop(heap)
        C[i] = CLOSED
        for inc in incs:
            j = i + inc
            if 0 <= j < n and C[j] == OPEN:
                if flat[j] < zmin:
                    flat[j] = zmin
                    n_raised += 1
                C[j] = ON_HEAP
                heapq.heappush(heap, (float(flat[j]), j))
    return DEM, n_raised


def satzone_dzw(d8, Q_gw, Rg, da, dt):
    """dzw of update_water_table before the layer bookkeeping
    (components/satzone_base.py:930-1008).  d8: dict of oracle.d8.build()."""
    dzw = dt * (Rg - Q_gw / da)
    if np.ndim(dzw) == 0:
        dzw = dzw + np.zeros_like(Q_gw)
    w, p = d8['w'], d8['p']
    if np.size(da) == 1:
        factor = (dt / da)
        for k in range(8):
            if p[k] is not None:
                dzw[p[k]] += (Q_gw[w[k]] * factor)
    else:
        for k in range(8):
            if p[k] is not None:
                dzw[p[k]] += (dt * Q_gw[w[k]] / da[p[k]])
    return dzw
