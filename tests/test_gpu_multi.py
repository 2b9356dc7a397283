"""Two row strips on two GPUs (one process each, halo rows over NCCL through
the C ABI) against one strip holding the whole grid: every owned row must be
bit-identical, for the D8 products and for the routed state."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NX, NY, STEPS = 200, 192, 60


def _n_devices():
    from topoflow_old_b200.engine import device_count
    return device_count()


def _inputs():
    from topoflow_old_b200.utils import synth
    dem = synth.fractal_dem(NX, NY, seed=5, roughness=8.0)
    dem[60:66, 40:52] = dem[60, 40]                 # a flat straddling nothing in particular
    dem[94:98, 100:110] = dem[95, 100]              # a flat straddling the strip boundary (row 96)
    w = np.float64(synth.uniform_field(NX, NY, 1.0, 6.0, seed=6))
    a = np.float64(synth.class_field(NX, NY, [0.0, 30.0, 45.0], seed=7, block=8)) * (np.pi / 180.0)
    n = np.float64(synth.uniform_field(NX, NY, 0.03, 0.15, seed=8))
    return dem, w, a, n


def _run_strip(rank, world, uid, wave, out_dir):
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200 import strips
    from topoflow_old_b200.engine import Strip
    from topoflow_old_b200.routing import setup_channels
    from topoflow_old_b200.utils import synth
    dem, w, a, n = _inputs()
    r0, nr = strips.strip_rows(NY, world, rank)
    s = Strip(NX, nr, device=rank if world > 1 else 0)
    if world > 1:
        north, south = strips.neighbours(world, rank)
        s.set_strip(r0, NY, north is not None, south is not None)
        s.comm_init(rank, world, uid)
    ones = np.ones(nr, dtype=np.float32)
    stats = s.d8_build(dem[r0:r0 + nr], 30 * ones, 30 * ones, np.float32(np.sqrt(1800.0)) * ones)
    code, ds = s.d8_get(L.D8_CODE), s.d8_get(L.D8_DS)
    # contributing area: with strips each rank accumulates its rows and the edge rows of the area
    # travel until no flow path has anything left to hand over (rows of different size: a
    # lat/lon-like da)
    da_rows = 900.0 * (1.0 + 0.001 * np.arange(r0, r0 + nr))
    n_unres = s.d8_area(da_rows, units_km=True)
    area = s.d8_get(L.D8_AREA)
    sl = lambda g: g[r0:r0 + nr].copy()
    slope = np.full((nr, NX), 0.01) + 0.001 * np.float64(sl(n))
    dt = 0.5 if wave == L.DYNAMIC else 2.0
    setup_channels(s, wave, L.MANNING, dt, 900.0, slope, sl(w), sl(a), sl(n),
                   outlets=[(nr // 2, NX // 2)], ds=ds)
    s.set_R(60.0 / 3.6e6)
    taps, _, bad = s.step(STEPS)
    assert bad[0] == -1 and bad[1] == -1
    out = dict(code=code, parent=s.d8_get(L.D8_PARENT_ID), inmask=s.d8_get(L.D8_INMASK),
               area=area, n_unres=np.array([n_unres]),
               vol=s.get(L.F_VOL), d=s.get(L.F_D), u=s.get(L.F_U), Q=s.get(L.F_Q),
               stats=np.array([stats['n_flats'], stats['n_pits'], stats['n_reps']]))
    np.savez(os.path.join(out_dir, 'w%d_r%d_wave%d.npz' % (world, rank, wave)), **out)
    s.close()


def _worker(rank, world, uid, wave, out_dir):
    _run_strip(rank, world, uid, wave, out_dir)


@pytest.mark.parametrize('wave', [0, 1, 2])
def test_two_gpu_strips_match_single_gpu(tmp_path, wave):
    if _n_devices() < 2:
        pytest.skip('needs 2 GPUs')
    import torch.multiprocessing as mp
    from topoflow_old_b200.engine import Strip
    uid = Strip.comm_unique_id()
    mp.spawn(_worker, args=(2, uid, wave, str(tmp_path)), nprocs=2, join=True)
    _run_strip(0, 1, None, wave, str(tmp_path))
    one = np.load(str(tmp_path / ('w1_r0_wave%d.npz' % wave)))
    parts = [np.load(str(tmp_path / ('w2_r%d_wave%d.npz' % (r, wave)))) for r in range(2)]
    for key in ('code', 'parent', 'inmask', 'area', 'vol', 'd', 'u', 'Q'):
        got = np.concatenate([p[key] for p in parts], axis=0)
        assert np.array_equal(got, one[key]), key
    assert one['area'].max() > 0.5 and one['area'].dtype == np.float32
    assert parts[0]['n_unres'][0] == parts[1]['n_unres'][0] == one['n_unres'][0]
    assert one['Q'].max() > 0
    assert np.array_equal(parts[0]['stats'], one['stats'])       # all-reduced counters


def _coupled_inputs():
    from topoflow_old_b200.utils import synth
    dem, w, a, n = _inputs()
    u = lambda seed: synth.uniform_field(NX, NY, 0.0, 1.0, seed=seed, dtype=np.float64)
    T_air = 6.0 - 12.0 * (np.float64(dem) - dem.min()) / (dem.max() - dem.min())
    met = dict(dt=30.0, P=(200.0 / 3.6e6) * (0.5 + u(51)), T_air=T_air, T_surf=T_air - 1.0, Qn_SW=300.0,
               Qn_LW=-50.0, rain_duration=100.0)
    snow = dict(dt=12.0, c0=2.7, T0=-0.2, h0_snow=0.3 * u(52), h0_swe=0.1 * u(52), rho_snow=300.0)
    evap = dict(dt=20.0, alpha=1.2, K_soil=0.45, soil_x=0.05, T_soil_x=0.0)
    infil = dict(dt=6.0, Ks=7.2e-6 * (0.5 + u(53)), Ki=1e-7, qs=0.485, qi=0.375808, G=0.724)
    return met, snow, evap, infil


def _run_coupled_strip(rank, world, uid, out_dir):
    """The coupled run of tests/test_gpu_coupled.py on strips: the providers work on their own
    rows, only the "did every cell change" test of the interpolation is global."""
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200 import strips
    from topoflow_old_b200.coupled import CoupledRun
    from topoflow_old_b200.engine import Strip
    from topoflow_old_b200.routing import setup_channels
    dem, w, a, n = _inputs()
    r0, nr = strips.strip_rows(NY, world, rank)
    s = Strip(NX, nr, device=rank if world > 1 else 0)
    if world > 1:
        north, south = strips.neighbours(world, rank)
        s.set_strip(r0, NY, north is not None, south is not None)
        s.comm_init(rank, world, uid)
    ones = np.ones(nr, dtype=np.float32)
    s.d8_build(dem[r0:r0 + nr], 30 * ones, 30 * ones, np.float32(np.sqrt(1800.0)) * ones)
    sl = lambda g: g[r0:r0 + nr].copy() if np.ndim(g) else g
    slope = np.full((nr, NX), 0.01) + 0.001 * np.float64(sl(n))
    setup_channels(s, L.KINEMATIC, L.MANNING, 2.0, 900.0, slope, sl(w), sl(a), sl(n),
                   outlets=[(nr // 2, NX // 2)], ds=s.d8_get(L.D8_DS))
    met, snow, evap, infil = [{k: sl(v) for k, v in d.items()} for d in _coupled_inputs()]
    run = CoupledRun(s, 2.0, met, snow, evap, infil)
    run.run_until(120.0)
    assert run.bad[0] == -1 and run.bad[1] == -1
    out = dict(vol=s.get(L.F_VOL), d=s.get(L.F_D), u=s.get(L.F_U), Q=s.get(L.F_Q), SM=run.get(run.SM.obj),
               ET=run.get(run.ET.obj), IN=run.get(run.IN.obj), I=run.get(run.I),
               lin=np.array([run.SM.linear, run.ET.linear, run.IN.linear]),
               vols=np.array([run.vol[k] for k in ('SM', 'ET', 'IN', 'R')]))
    np.savez(os.path.join(out_dir, 'c%d_r%d.npz' % (world, rank)), **out)
    s.close()


def _coupled_worker(rank, world, uid, out_dir):
    _run_coupled_strip(rank, world, uid, out_dir)


def test_two_gpu_coupled_run_matches_single_gpu(tmp_path):
    if _n_devices() < 2:
        pytest.skip('needs 2 GPUs')
    import torch.multiprocessing as mp
    from topoflow_old_b200.engine import Strip
    uid = Strip.comm_unique_id()
    mp.spawn(_coupled_worker, args=(2, uid, str(tmp_path)), nprocs=2, join=True)
    _run_coupled_strip(0, 1, None, str(tmp_path))
    one = np.load(str(tmp_path / 'c1_r0.npz'))
    parts = [np.load(str(tmp_path / ('c2_r%d.npz' % r))) for r in range(2)]
    for key in ('vol', 'd', 'u', 'Q', 'SM', 'ET', 'IN', 'I'):
        got = np.concatenate([p[key] for p in parts], axis=0)
        assert np.array_equal(got, one[key]), key
    for p in parts:
        assert np.array_equal(p['lin'], one['lin'])              # the same interpolation decisions
    vols = parts[0]['vols'] + parts[1]['vols']
    assert np.allclose(vols, one['vols'], rtol=1e-11, atol=0)
    assert one['Q'].max() > 0
