"""Pin the D8 oracle against the reference's own golden grids and tables."""
import os

import numpy as np
import pytest

from oracle import d8 as od8

# RT3 resolve array, topoflow/components/tests/test_d8_base_UNFINISHED.py:101-125
RT3_RESOLVE = """
  0   1   2   2   4   1   2   2   8   8   2   2   8   8   2   2
 16   1   2   2   4   4   2   2   8   8   8   2   8   8   8   2
 32  32   2   2  32  32   2   2   8   8   8   2   8   8   8   2
 32  32  32  32  32  32   2   2   8   8   8   8   8   8   8   8
 64  64   2   2   4   1   2   2   8   8   2   2   8   8   2   2
 16  64   2   2  16   1   2   2   8   8   8   2   8   8   8   2
 32  32  32  32  32  32   2   2  32  32  32   2   8   8   8   2
 32  32  32  32  32  32  32   2  32  32  32  32   8   8   8   8
128 128 128 128 128 128   2   2   8 128   2   2   8 128   2   2
128 128 128 128 128 128   2   2   8   8   8   2   8   8   8   2
 32 128 128 128  32 128   2   2  32 128   2   2   8   8   8   2
 32  32  32 128  32  32  32   2  32  32  32   8   8   8   8   2
128 128 128 128 128 128 128 128 128 128 128 128 128 128   2   2
128 128 128 128 128 128 128 128   8 128 128 128   8   8   8   2
 32 128 128 128  32 128 128 128  32 128 128 128  32 128   2 128
 32  32  32 128  32  32  32 128  32  32  32  32  32  32   8   8
"""


def treynor(golden_dir, name, dtype):
    a = np.fromfile(os.path.join(golden_dir, 'treynor', name), dtype=dtype)
    return a.reshape(44, 29)


def test_resolve_table_matches_rt3():
    t = od8.resolve_table()
    want = np.array(RT3_RESOLVE.split(), dtype=np.int64)
    assert t.sum() == 10161
    assert np.array_equal(t.astype(np.int64), want)
    # single directions map to themselves (d8_global.py:356-359)
    for g in od8.CODE_LIST:
        assert t[g] == g


def test_treynor_flow_codes_bit_exact(golden_dir):
    dem = treynor(golden_dir, 'Treynor_DEM.rtg', '>f4').astype(np.float32)
    want = treynor(golden_dir, 'Treynor_flow.rtg', 'u1')
    dx, dy, dd = od8.pixel_sizes(44, 30.0, 30.0)
    got, stats = od8.flow_grid(dem, dx[0], dy[0], dd[0], return_stats=True)
    assert got.dtype == np.uint8
    assert np.array_equal(got, want)
    # counters printed in Test1_kin_wave/00_Full_Output.txt
    assert stats['n_flats'] == 2 and stats['n_pits'] == 5
    assert stats['n_reps'] == 1 and stats['n_bad'] == 0


def test_treynor_area_bit_exact(golden_dir):
    code = treynor(golden_dir, 'Treynor_flow.rtg', 'u1')
    want = treynor(golden_dir, 'Treynor_area.rtg', '>f4').astype(np.float32)
    A, reps = od8.area_grid(code, np.float64(30.0) * np.float64(30.0),
                            'km^2', return_reps=True)
    assert A.dtype == np.float32
    assert np.array_equal(A.view(np.uint32), want.view(np.uint32))
    assert np.float32(A.max()) == np.float32(0.49860007)
    assert np.array_equal(od8.area_grid_fast(code, 900.0).view(np.uint32),
                          want.view(np.uint32))


def test_treynor_ds_dw_and_parents(golden_dir):
    dem = treynor(golden_dir, 'Treynor_DEM.rtg', '>f4').astype(np.float32)
    b = od8.build(dem, 30.0, 30.0)
    # "min(dw), max(dw) = 30.0, 42.4264" in the golden log
    for g in (b['ds'], b['dw']):
        assert g.dtype == np.float32
        assert g.min() == np.float32(30.0)
        assert g.max() == np.float32(np.sqrt(1800.0))
    pid = b['parent_ID_grid']
    assert pid.dtype == np.int32
    d8 = b['d8_grid']
    assert np.all(pid[d8 == 0] == 0)
    # each w_k / p_k pair is one step in direction k
    for k in range(8):
        (wr, wc), p = b['w'][k], b['p'][k]
        if p is None:
            assert wr.size == 0
            continue
        dr, dc = od8.NBR[k]
        assert np.array_equal(p[0], wr + dr) and np.array_equal(p[1], wc + dc)
    assert np.array_equal(np.sort(np.ravel_multi_index(b['noflow_IDs'], d8.shape)),
                          np.where(d8.ravel() == 0)[0])


@pytest.mark.parametrize('seed,shape', [(1, (37, 53)), (2, (64, 64)), (3, (5, 7))])
def test_area_fast_equals_level_loop(seed, shape):
    rng = np.random.default_rng(seed)
    dem = (rng.random(shape) * 50).astype(np.float32)
    dem[rng.random(shape) < 0.05] = np.float32(10.0)      # flats
    dx, dy, dd = od8.pixel_sizes(shape[0], 30.0, 20.0)
    code = od8.flow_grid(dem, dx[0], dy[0], dd[0])
    a1 = od8.area_grid(code, 600.0)
    a2 = od8.area_grid_fast(code, 600.0)
    assert np.array_equal(a1.view(np.uint32), a2.view(np.uint32))
    da = (rng.random(shape) * 1000 + 1).astype(np.float64)
    assert np.array_equal(od8.area_grid(code, da, 'm^2').view(np.uint32),
                          od8.area_grid_fast(code, da, 'm^2').view(np.uint32))


def test_flats_pits_nodata_and_periodic():
    dem = np.full((8, 9), 5.0, dtype=np.float32)          # one big flat
    dem[3, 4] = 1.0                                        # a pit
    dem[5, 2] = np.nan
    dem[2, 6] = -9999.0
    dx, dy, dd = od8.pixel_sizes(8, 10.0, 10.0)
    raw, n_flats, n_pits, n_bad = od8.start_codes(dem, dx[0], dy[0], dd[0])
    assert raw[3, 4] == -300 and n_pits >= 1 and n_bad == 2
    assert raw[5, 2] == 0 and raw[2, 6] == 0
    assert np.all(raw[0] == 0) and np.all(raw[:, 0] == 0)
    code = od8.flow_grid(dem, dx[0], dy[0], dd[0])
    assert set(np.unique(code)) <= {0, 1, 2, 4, 8, 16, 32, 64, 128}
    # cells next to the nodata hole drain into it; the pit stays unresolved
    assert code[3, 4] == 0
    # periodic: borders keep their codes and parents wrap
    rng = np.random.default_rng(0)
    dem = (rng.random((6, 7)) * 9).astype(np.float32)
    c = od8.flow_grid(dem, dx[0], dy[0], dd[0], lr_periodic=True, tb_periodic=True)
    pid = od8.parent_id_grid(c, True, True)
    assert pid.min() >= 0 and pid.max() < 42
    r, q = np.divmod(np.arange(42), 7)
    for i in np.where(c.ravel() > 0)[0]:
        k = int(np.log2(c.ravel()[i]))
        dr, dc = od8.NBR[k]
        assert pid.ravel()[i] == ((r[i] + dr) % 6) * 7 + (q[i] + dc) % 7
