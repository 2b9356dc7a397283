"""NumPy oracle for the D8 flow-grid builder (TEST INFRASTRUCTURE ONLY).

Restates, whole-array operation by whole-array operation, what the reference
does in topoflow/components/d8_base.py and d8_global.py so that its results
are bit-identical to the reference's (D8 codes: uint8; area / ds / dw:
float32; parent IDs: int32).  Every function cites the reference lines it
follows.  Legacy-NumPy behaviour the reference relied on (value-based casting,
capitalised dtype names, Python-2 integer division) is made explicit here.

Parity: pinned against Treynor_flow.rtg / Treynor_area.rtg and the RT3
tie-break table (tests/test_oracle_d8.py).

Neighbour subscripts k = 0..7 are UR, R, LR, B, LL, L, UL, T
(d8_global.py:381-385); flow code g[k] = 1 << k points at neighbour k.
Row index grows southward (row 0 = north).
"""
import numpy as np

# d8_base.py:616-633 (Jenson-84 / RiverTools codes) and :637-648 (opposites)
CODE_LIST = np.array([1, 2, 4, 8, 16, 32, 64, 128], dtype=np.int16)
CODE_OPPS = np.array([16, 32, 64, 128, 1, 2, 4, 8], dtype=np.int16)
# (d_row, d_col) of neighbour k; d8_global.py:225-232 / :458-465
NBR = ((-1, 1), (0, 1), (1, 1), (1, 0), (1, -1), (0, -1), (-1, -1), (-1, 0))
NODATA = np.float32(-9999)       # d8_base.py:184
PIT_CODE = -300                  # d8_global.py:302


# ----------------------------------------------------------------------------
# static tables
# ----------------------------------------------------------------------------
def _rotate_codes(v, times):
    """Rotate byte-coded direction sets by 45 deg `times` times
    (d8_base.py:744-762: double, wrap the carry bit, mask to 8 bits)."""
    v = np.asarray(v, dtype=np.int64)
    for _ in range(times):
        v = v * 2
        v = v + (v > 255)
        v = v & 255
    return v


def resolve_table():
    """256-entry tie-break table `resolve` (d8_base.py:766-816)."""
    t = np.zeros(256, dtype=np.uint8)
    # high-symmetry groups (:787-792)
    for value, members in ((1, (17, 85)),
                           (2, (34, 102, 119, 238, 170)),
                           (8, (136, 153, 187, 221, 255)),
                           (4, (68,)), (32, (51,)), (128, (204,))):
        t[list(members)] = value
    # corner cases (:797-800)
    for value, members in ((1, (1, 5, 69)), (4, (4, 20, 21)),
                           (16, (16, 80, 84)), (64, (64, 65, 81))):
        t[list(members)] = value
    # rotationally equivalent families (:802-814)
    fam = np.array([2, 3, 6, 7, 10, 11, 14, 15, 18, 19, 22, 23,
                    27, 31, 35, 38, 39, 43, 47, 54, 55, 66, 67, 70,
                    71, 74, 75, 78, 79, 82, 83, 86, 87, 91, 95, 103,
                    107, 111, 134, 135, 138, 139, 142, 143, 150, 151, 155,
                    159, 166, 167, 171, 175, 183, 191, 206, 207, 223],
                   dtype=np.int64)
    t[fam] = 2
    t[_rotate_codes(fam, 2)] = 8
    t[_rotate_codes(fam, 4)] = 32
    t[_rotate_codes(fam, 6)] = 128
    return t


def valid_code_map():
    """uint8[256]: the 8 legal codes map to themselves, all else to 0
    (d8_base.py:652-660)."""
    m = np.zeros(256, dtype=np.uint8)
    m[CODE_LIST] = CODE_LIST.astype(np.uint8)
    return m


def inc_map(nx):
    """int32[129]: parent_ID - ID for each flow code (d8_base.py:679-693)."""
    incs = np.array([-nx + 1, 1, nx + 1, nx, nx - 1, -1, -nx - 1, -nx],
                    dtype=np.int32)
    m = np.zeros(129, dtype=np.int32)
    m[CODE_LIST] = incs
    return m


def edge_ids(nx, ny):
    """(rows, cols) of border cells, corners once (d8_base.py:697-718)."""
    top = np.arange(nx, dtype=np.int32)
    bot = top + (ny - 1) * nx
    left = (1 + np.arange(ny - 2, dtype=np.int32)) * nx
    right = left + (nx - 1)
    ids = np.concatenate([top, bot, left, right])
    return divmod(ids, nx)


def not_edge_grid(nx, ny):
    """uint8 grid, 0 on the border (d8_base.py:728-735)."""
    g = np.ones((ny, nx), dtype=np.uint8)
    g[edge_ids(nx, ny)] = 0
    return g


def pixel_sizes(ny, xres, yres):
    """dx, dy, dd per row as float32 for fixed-length pixels
    (pixels.py:136-148 with METERS=True, then d8_base.py:565-573)."""
    dx = np.float64(xres)
    dy = np.float64(yres)
    dd = np.sqrt(dx ** 2 + dy ** 2)
    z = np.zeros(ny, dtype=np.float64)
    return (np.float32(z + dx), np.float32(z + dy), np.float32(z + dd))


# ----------------------------------------------------------------------------
# flow codes
# ----------------------------------------------------------------------------
def _shifted(a, k):
    """Grid whose (r, c) entry is a[(r+dr) % ny, (c+dc) % nx] for neighbour k
    (np.roll, periodic: d8_global.py:225-232)."""
    dr, dc = NBR[k]
    return np.roll(np.roll(a, -dr, axis=0), -dc, axis=1)


def start_codes(dem, dx0, dy0, dd0, link_flats=True,
                lr_periodic=False, tb_periodic=False, nodata=NODATA):
    """int16 raw codes with ties OR-ed (d8_global.py:179-334).

    Returns (codes, n_flats, n_pits, n_bad).  Flats (max slope == 0) carry
    minus the OR of their tied directions, pits -300, nodata / NaN cells and
    (non-periodic) borders 0."""
    dem = np.asarray(dem)
    ny, nx = dem.shape
    dist = (dd0, dx0, dd0, dy0, dd0, dx0, dd0, dy0)      # :225-232
    slopes = [(dem - _shifted(dem, k)) / dist[k] for k in range(8)]
    smax = np.maximum(slopes[0], slopes[1])              # :250-256 (NaN propagates)
    for k in range(2, 8):
        np.maximum(smax, slopes[k], smax)
    codes = (slopes[0] == smax) * CODE_LIST[0]           # :266-274 -> int16
    for k in range(1, 8):
        codes = codes + (slopes[k] == smax) * CODE_LIST[k]
    codes = codes.astype(np.int16)
    n_flats = n_pits = 0
    if not link_flats:                                   # :281-285
        codes[smax <= 0] = 0
    else:                                                # :286-302
        flats = (smax == 0)
        n_flats = int(flats.sum())
        codes[flats] = -1 * codes[flats]
        pits = (smax < 0)
        n_pits = int(pits.sum())
        codes[pits] = PIT_CODE
    bad = np.logical_or(dem <= nodata, np.isfinite(dem) != 1)   # :312-315
    n_bad = int(bad.sum())
    codes[bad] = 0
    if not lr_periodic:                                  # :320-322
        codes[:, 0] = 0
        codes[:, nx - 1] = 0
    if not tb_periodic:                                  # :327-329
        codes[0, :] = 0
        codes[ny - 1, :] = 0
    return codes, n_flats, n_pits, n_bad


def break_ties(codes, resolve):
    """codes > 0 -> resolve[code] in place (d8_global.py:352-367)."""
    w = codes > 0
    codes[w] = resolve[codes[w]]
    return codes


def link_flats(codes, resolve, not_edge=None):
    """Iteratively give each flat a direction toward an already-resolved
    neighbour that does not flow back (d8_global.py:371-567).  Jacobi: all
    neighbour codes are read before any flat is rewritten in an iteration.
    Returns (codes, n_reps, total_fixed)."""
    ny, nx = codes.shape
    if not_edge is None:
        not_edge = not_edge_grid(nx, ny)
    g, h = CODE_LIST, CODE_OPPS
    total = 0
    n_reps = 0
    while True:
        n_reps += 1
        rows, cols = np.where((codes < 0) & (codes != PIT_CODE) & (not_edge == 1))
        n_flats = rows.size
        if n_flats == 0:
            break
        ready = np.zeros(n_flats, dtype=np.int16)
        left = -1 * codes[rows, cols]                    # :505
        allowed = [None] * 8
        for k in range(7, 0, -1):                        # :507-520, high bit first
            allowed[k] = left >= g[k]
            left = left - g[k] * allowed[k]
        allowed[0] = left >= g[0]
        for k in range(8):
            dr, dc = NBR[k]
            nb = codes[(rows + dr) % ny, (cols + dc) % nx]       # :445-465
            valid = (nb > 0) & (nb != h[k])                      # :491-498
            ready = ready + (valid & allowed[k]) * g[k]          # :529-536
        fix = np.where(ready > 0)[0]
        if fix.size:
            total += fix.size
            codes[rows[fix], cols[fix]] = resolve[ready[fix]]    # :546-548
        if fix.size == n_flats or fix.size == 0:                 # :555-557
            break
    return codes, n_reps, total


def finalize_codes(codes, vmap=None):
    """int16 working codes -> uint8 D8 grid (d8_global.py:141-144)."""
    if vmap is None:
        vmap = valid_code_map()
    c = codes.copy()
    c[(c < 0) | (c > 128)] = 0
    return vmap[c]


def flow_grid(dem, dx0, dy0, dd0, link=True, lr_periodic=False,
              tb_periodic=False, return_stats=False):
    """update_flow_grid() (d8_global.py:97-144) with BREAK_TIES on."""
    resolve = resolve_table()
    codes, n_flats, n_pits, n_bad = start_codes(
        dem, dx0, dy0, dd0, link_flats=link,
        lr_periodic=lr_periodic, tb_periodic=tb_periodic)
    break_ties(codes, resolve)
    n_reps = 0
    if link:
        ny, nx = codes.shape
        codes, n_reps, _ = link_flats(codes, resolve, not_edge_grid(nx, ny))
    d8 = finalize_codes(codes)
    if return_stats:
        return d8, dict(n_flats=n_flats, n_pits=n_pits, n_bad=n_bad,
                        n_reps=n_reps)
    return d8


# ----------------------------------------------------------------------------
# derived index structures
# ----------------------------------------------------------------------------
def parent_id_grid(d8, lr_periodic=False, tb_periodic=False):
    """int32 grid of downstream-cell IDs; 0 for no-flow cells and
    (non-periodic) borders (d8_global.py:574-644)."""
    ny, nx = d8.shape
    n_pixels = nx * ny
    ids = np.arange(n_pixels, dtype=np.int32).reshape(ny, nx)   # d8_base.py:675
    pid = ids + inc_map(nx)[d8]
    dirs = d8.reshape(-1)
    flat = pid.reshape(-1)
    top = np.arange(nx, dtype=np.int32)
    bot = top + nx * (ny - 1)
    lft = nx * np.arange(ny, dtype=np.int32)
    rgt = lft + (nx - 1)
    if lr_periodic:                                             # :609-616
        w = np.isin(dirs[rgt], (1, 2, 4))
        flat[rgt[w]] -= nx
        w = np.isin(dirs[lft], (16, 32, 64))
        flat[lft[w]] += nx
    else:
        flat[rgt] = 0
        flat[lft] = 0
    if tb_periodic:                                             # :625-632
        w = np.isin(dirs[top], (1, 64, 128))
        flat[top[w]] += n_pixels
        w = np.isin(dirs[bot], (4, 8, 16))
        flat[bot[w]] -= n_pixels
    else:
        flat[top] = 0
        flat[bot] = 0
    pid[d8 <= 0] = 0                                            # :641-644
    return pid


def parent_ids(pid):
    """(rows, cols) of every cell's parent (d8_global.py:648-675)."""
    return divmod(pid, pid.shape[1])


def flow_from_to(d8, pid):
    """w_k = where(code == g_k) in row-major order, p_k = their parents
    (d8_global.py:719-835).  Returns lists w[0..7], p[0..7] (p[k] None when
    direction k does not occur)."""
    nx = d8.shape[1]
    w, p = [], []
    for k in range(8):
        wk = np.where(d8 == CODE_LIST[k])
        w.append(wk)
        p.append(divmod(pid[wk], nx) if wk[0].size else None)
    return w, p


def noflow_ids(d8):
    """Cells where flow terminates (d8_global.py:839-859)."""
    w = np.where(d8 <= 0)
    if w[0].size == 0:
        return edge_ids(d8.shape[1], d8.shape[0])
    return w


def flow_width_grid(d8, dx, dy, dd):
    """float32 dw: diagonal -> dd[row], E/W -> dy[row], N/S and undefined ->
    dx[row] (d8_global.py:863-935, METHOD2 off)."""
    ny, nx = d8.shape
    dw = np.zeros((ny, nx), dtype=np.float32)
    rows = np.broadcast_to(np.arange(ny)[:, None], (ny, nx))
    for codes, src in (((1, 4, 16, 64), dd), ((2, 32), dy),
                       ((8, 128), dx), ((0,), dx)):
        m = np.isin(d8, codes)
        dw[m] = src[rows[m]]
    return dw


def flow_length_grid(d8, dx, dy, dd):
    """float32 ds: diagonal -> dd[row], E/W -> dx[row], N/S -> dy[row],
    undefined -> dx[row] (d8_global.py:949-1004)."""
    ny, nx = d8.shape
    ds = np.zeros((ny, nx), dtype=np.float32)
    rows = np.broadcast_to(np.arange(ny)[:, None], (ny, nx))
    for codes, src in (((1, 4, 16, 64), dd), ((2, 32), dx),
                       ((8, 128), dy), ((0,), dx)):
        m = np.isin(d8, codes)
        ds[m] = src[rows[m]]
    return ds


# ----------------------------------------------------------------------------
# contributing area
# ----------------------------------------------------------------------------
def area_grid(d8, da, a_units='km^2', return_reps=False):
    """float32 contributing area, the reference's own level-by-level loop
    (d8_global.py:1018-1261).  `da` is a scalar [m^2] or an [ny, nx] grid.
    Each iteration reads the pre-iteration A of all 8 neighbours, then every
    READY cell gets pixel_area followed by += child areas for k = 0..7."""
    ny, nx = d8.shape
    A = np.zeros((ny, nx), dtype=np.float32)
    h = CODE_OPPS
    pixel_area = np.asarray(da, dtype=np.float64)
    if 'km' in a_units.lower():                                 # :1064-1067
        pixel_area = pixel_area / 1e6
    n_reps = 0
    while True:
        n_reps += 1
        rows, cols = np.where((A == 0) & (d8 != 0))             # :1083
        if rows.size == 0:
            break
        a_nb, is_kid = [], []
        for k in range(8):
            dr, dc = NBR[k]
            rr, cc = (rows + dr) % ny, (cols + dc) % nx         # :1105-1134
            a_nb.append(A[rr, cc])
            is_kid.append(d8[rr, cc] == h[k])
        ready = np.ones(rows.size, dtype=bool)
        for k in range(8):                                      # :1161-1191
            ready &= ~(is_kid[k] & (a_nb[k] == 0))
        wr = np.where(ready)[0]
        if wr.size:
            r, c = rows[wr], cols[wr]
            if pixel_area.size == 1:                            # :1209-1212
                A[r, c] = pixel_area
            else:
                A[r, c] = pixel_area[r, c]
            for k in range(8):                                  # :1214-1244
                kk = np.where(is_kid[k][wr])[0]
                if kk.size:
                    A[r[kk], c[kk]] += a_nb[k][wr[kk]]
        if wr.size == rows.size or wr.size == 0:                # :1249-1258
            break
    if return_reps:
        return A, n_reps
    return A


def area_grid_fast(d8, da, a_units='km^2'):
    """Same float32 result as area_grid() in O(N): Kahn levels over the flow
    tree, each cell summing pixel_area then its children in neighbour order
    k = 0..7 in float32.  Valid because a cell's value is written once, after
    all of its children are final (d8_global.py:1161-1244), so the result
    does not depend on which iteration wrote it.  Checked against area_grid()
    in tests/test_oracle_d8.py; used only to reach sizes the level loop
    cannot."""
    ny, nx = d8.shape
    n = nx * ny
    pixel_area = np.asarray(da, dtype=np.float64)
    if 'km' in a_units.lower():
        pixel_area = pixel_area / 1e6
    code = d8.reshape(-1)
    rr, cc = np.divmod(np.arange(n, dtype=np.int64), nx)
    # parent (periodic wrap, as the % in the reference) of every flowing cell
    kidx = np.zeros(256, dtype=np.int64)
    kidx[CODE_LIST] = np.arange(8)
    k_of = kidx[code]
    drs = np.array([o[0] for o in NBR])[k_of]
    dcs = np.array([o[1] for o in NBR])[k_of]
    parent = ((rr + drs) % ny) * nx + ((cc + dcs) % nx)
    flowing = code != 0
    parent[~flowing] = -1
    # a parent with code 0 never becomes "unknown", so it never sums kids
    ok = flowing.copy()
    ok[flowing] &= flowing[parent[flowing]]
    pending = np.bincount(parent[ok], minlength=n)
    A = np.zeros(n, dtype=np.float32)
    frontier = np.where(flowing & (pending == 0))[0]
    # child slot table: kid[k][i] = id of the child of i at neighbour k or -1
    kid = np.full((8, n), -1, dtype=np.int64)
    src = np.where(ok)[0]
    # child at neighbour k of its parent has code h[k]; slot = opposite index
    slot = (k_of[src] + 4) % 8
    kid[slot, parent[src]] = src
    while frontier.size:
        if pixel_area.size == 1:
            acc = np.full(frontier.size, np.float32(pixel_area), dtype=np.float32)
        else:
            acc = pixel_area.reshape(-1)[frontier].astype(np.float32)
        for k in range(8):
            ck = kid[k, frontier]
            m = ck >= 0
            acc[m] = acc[m] + A[ck[m]]
        A[frontier] = acc
        par = parent[frontier]
        par = par[ok[frontier]]
        np.subtract.at(pending, par, 1)
        cand = np.unique(par)
        frontier = cand[pending[cand] == 0]
    return A.reshape(ny, nx)


def build(dem, xres, yres, link=True, lr_periodic=False, tb_periodic=False,
          with_area=False, a_units='km^2'):
    """d8.update_most() / update() (d8_base.py:351-479) as one call.
    Returns a dict with the fields the channels component reads."""
    ny, nx = dem.shape
    dx, dy, dd = pixel_sizes(ny, xres, yres)
    d8, stats = flow_grid(dem, dx[0], dy[0], dd[0], link=link,
                          lr_periodic=lr_periodic, tb_periodic=tb_periodic,
                          return_stats=True)
    pid = parent_id_grid(d8, lr_periodic, tb_periodic)
    out = dict(d8_grid=d8, parent_ID_grid=pid, parent_IDs=parent_ids(pid),
               dw=flow_width_grid(d8, dx, dy, dd),
               ds=flow_length_grid(d8, dx, dy, dd),
               noflow_IDs=noflow_ids(d8), dx=dx, dy=dy, dd=dd, stats=stats)
    out['w'], out['p'] = flow_from_to(d8, pid)
    if with_area:
        da = np.float64(xres) * np.float64(yres)      # pixels.py:174-185
        out['A'] = area_grid(d8, da, a_units)
    return out
