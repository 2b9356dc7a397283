"""The drop-in BMI components driven the way EMELI drives them: from the
reference's own cfg / RTI / RTG files, through initialize / update / finalize,
get_values / set_values.  Results are compared with the reference's shipped
output files and with the oracle."""
import json
import os
import shutil

import numpy as np
import pytest

from oracle import channels as och
from oracle import d8 as od8
from tests.test_gpu_channels import RTOL, field_err

pytestmark = pytest.mark.gpu

ZERO_INPUTS = ('snowpack__melt_volume_flux', 'land_surface_water__baseflow_volume_flux',
               'glacier_ice__melt_volume_flux', 'land_surface_water__evaporation_volume_flux',
               'soil_surface_water__infiltration_volume_flux')


def stage(golden_dir, case, tmp_path):
    dst = str(tmp_path / case)
    shutil.copytree(os.path.join(golden_dir, case), dst)
    for f in os.listdir(dst):
        if f.endswith('_path_info.cfg'):
            p = os.path.join(dst, f)
            txt = open(p).read().replace('~/TF_Output/Treynor', '.')
            open(p, 'w').write(txt)
    return dst


def connect_disabled_providers(comp, P_rain):
    """What EMELI's initialize_comp_set does for the channels component when
    every other process is Disabled: 0-d zeros by reference."""
    comp.set_values('atmosphere_water__rainfall_volume_flux', P_rain)
    for name in ZERO_INPUTS:
        comp.set_values(name, och.scalar(0))
    comp.set_values('water-liquid__mass-per-volume_density', och.scalar(1000.0))
    for name in ('canals__count', 'sinks__count', 'sources__count'):
        comp.set_values(name, np.array(0, dtype='int32'))


def test_plane_case_reproduces_the_reference_output_files(golden_dir, tmp_path):
    from topoflow_old_b200.components import channels_kinematic_wave
    d = stage(golden_dir, 'plane_canal', tmp_path)
    rain = np.fromfile(os.path.join(d, 'Rain_100mmph_on_right.rtg'), dtype='<f4').reshape(20, 20)
    P = np.float64(rain * np.float32(np.float64(1) / np.float64(3600000)))
    comp = channels_kinematic_wave.channels_component()
    assert comp.get_status() == 'created'
    comp.initialize(cfg_file=os.path.join(d, 'Case1_channels_kinematic_wave.cfg'), SILENT=True)
    assert comp.get_status() == 'initialized'
    # everything is scalar in this cfg: width 10, angle 0, n 0.3, dt 6
    assert float(comp.dt) == 6.0 and comp.width.ndim == 0 and comp.n_outlets == 12
    connect_disabled_providers(comp, P)
    Q_outlet = comp.get_values('basin_outlet_water_x-section__volume_flow_rate')
    Q_grid = comp.get_values('channel_water_x-section__volume_flow_rate')
    for n in range(600):
        comp.update(-1.0)
        assert comp.get_status() == 'updated'
    # references handed out before the run see the final state
    assert Q_outlet is comp.Q_outlet and abs(float(Q_outlet) - 0.025) < 1e-6
    assert Q_grid is comp.get_values('channel_water_x-section__volume_flow_rate')
    assert abs(Q_grid[1, 18] - 0.05) < 2e-6
    assert float(comp.get_current_time()) == 3600.0
    comp.finalize()
    assert comp.get_status() == 'finalized'
    # ---- text hydrographs: Q, d, u at 12 cells every 30 s ----
    for var in 'Qdu':
        mine = open(os.path.join(d, 'Case1_0D-%s.txt' % var)).read().splitlines()
        gold = open(os.path.join(d, 'out', 'Case1_0D-%s.txt' % var)).read().splitlines()
        assert mine[:2] == gold[:2] and len(mine) == len(gold) == 122
        a = np.array([r.split() for r in mine[2:]], dtype=float)
        b = np.array([r.split() for r in gold[2:]], dtype=float)
        assert np.array_equal(a[:, 0], b[:, 0])
        # columns (10,18) and (1,18) are independent of the canal the golden run had
        assert np.max(np.abs(a[:, 1:3] - b[:, 1:3])) <= 1.0000001e-7, var
    # ---- grid stack: 60 frames of float32 Q, same bytes/frame and layout ----
    from topoflow_old_b200.utils import rts_files
    mine, gold = rts_files.rts_file(), rts_files.rts_file()
    assert mine.open_file(os.path.join(d, 'Case1_2D-Q.rts'))
    assert gold.open_file(os.path.join(d, 'out', 'Case1_2D-Q.rts'))
    assert mine.number_of_grids() == gold.number_of_grids() == 60
    cols = [c for c in range(20) if c not in (3, 16)]        # canal exit / entrance columns
    for k in (0, 1, 10, 30, 59):
        a, b = mine.get_grid(k)[:, cols], gold.get_grid(k)[:, cols]
        assert np.max(np.abs(a - b)) <= 1e-6 * max(1e-30, np.abs(b).max())
    assert os.path.exists(os.path.join(d, 'Case1_2D-Q.rti')) and os.path.exists(os.path.join(d, 'Case1_2D-Q.bov'))


def test_plane_canal_run_through_the_component(golden_dir, tmp_path):
    """The shipped Case1 run as it was made: the CUDA component in the channels
    seat of the replayed EMELI loop, with the canal.  All columns of the three
    text files and every frame of the Q stack."""
    from tests import replay
    from topoflow_old_b200.components import channels_kinematic_wave
    d = stage(golden_dir, 'plane_canal', tmp_path)
    comp = channels_kinematic_wave.channels_component()
    cfg = os.path.join(d, 'Case1_channels_kinematic_wave.cfg')
    fw, chan = replay.run_plane(golden_dir, lambda div: replay.BmiPlaneChannels(comp, cfg, div))
    assert int(comp.n_canals) == 1 and comp.canal_in_IDs[0][0] == 10 and comp.canal_in_IDs[1][0] == 16
    assert comp.canal_out_IDs[1][0] == 3
    comp.finalize()
    for var in 'Qdu':
        mine = open(os.path.join(d, 'Case1_0D-%s.txt' % var)).read().splitlines()
        gold = open(os.path.join(d, 'out', 'Case1_0D-%s.txt' % var)).read().splitlines()
        a = np.array([r.split() for r in mine[2:122]], dtype=float)
        b = np.array([r.split() for r in gold[2:]], dtype=float)
        assert a.shape == b.shape == (120, 13)
        assert np.max(np.abs(a - b)) <= 1.0000001e-7, var      # one unit of the 7th printed decimal
        assert np.mean(a == b) > 0.999
    mine = np.fromfile(os.path.join(d, 'Case1_2D-Q.rts'), dtype='<f4')[:60 * 400].reshape(60, 20, 20)
    gold = np.fromfile(os.path.join(d, 'out', 'Case1_2D-Q.rts'), dtype='<f4').reshape(60, 20, 20)
    ulp = np.abs(mine.view(np.int32).astype(np.int64) - gold.view(np.int32).astype(np.int64))
    assert ulp.max() <= 1 and np.mean(ulp == 0) > 0.999         # fp64 within 1e-10 -> float32 rounding ties
    assert gold[:, 10, 3].max() > 0.02 and mine[:, 10, 16].max() == 0.0


@pytest.mark.parametrize('wave', ['kinematic', 'diffusive', 'dynamic'])
def test_treynor_component_matches_oracle(golden_dir, tmp_path, wave):
    import importlib
    mod = importlib.import_module('topoflow_old_b200.components.channels_%s_wave' % wave)
    d = stage(golden_dir, 'treynor', tmp_path)
    comp = mod.channels_component()
    comp.initialize(cfg_file=os.path.join(d, 'June_20_67_channels_%s_wave.cfg' % wave), SILENT=True)
    log = json.load(open(os.path.join(golden_dir, 'treynor', 'log_scalars.json')))['Test1_kin_wave']
    # D8 grid built on the GPU from Treynor_DEM.rtg == the shipped flow grid
    flow = np.fromfile(os.path.join(d, 'Treynor_flow.rtg'), dtype='u1').reshape(44, 29)
    assert np.array_equal(comp.d8.d8_grid, flow)
    assert comp.d8.d8_stats['n_flats'] == log['n_flats'] and comp.d8.d8_stats['n_pits'] == log['n_pits']
    assert comp.outlet_ID == (34, 16)
    # the lazily built index lists agree with the reference's definitions
    b = od8.build(np.float32(comp.d8.DEM), 30.0, 30.0)
    assert np.array_equal(comp.d8.w3[0], b['w'][2][0]) and np.array_equal(comp.d8.p3[1], b['p'][2][1])
    assert np.array_equal(comp.d8.noflow_IDs[0], b['noflow_IDs'][0])
    assert np.array_equal(comp.d8.parent_IDs[0], b['parent_IDs'][0])
    # same inputs into the oracle
    rd = lambda f: np.fromfile(os.path.join(d, f), dtype='>f4').reshape(44, 29).astype(np.float32)
    inp = och.prepare_inputs(b, rd('Treynor_slope.rtg'), rd('Treynor_chan-w.rtg'),
                             rd('Treynor_chan-a.rtg'), nval=rd('Treynor_chan-n.rtg'))
    ch = och.ChannelsOracle(b, {'kinematic': 0, 'diffusive': 1, 'dynamic': 2}[wave], dt=float(comp.dt),
                            da=np.float64(900.0), slope=inp['slope'], width=inp['width'],
                            angle=inp['angle'], nval=inp['nval'], outlet_ID=(34, 16))
    P = och.scalar(0.0)
    ch.P_rain = P
    connect_disabled_providers(comp, P)
    for n in range(400):
        P.fill(80.0 / 3.6e6 if n < 250 else 0.0)     # a rain pulse, seen by reference
        comp.update(-1.0)
        ch.update()
    for long_name, attr in (('channel_water_x-section__volume_flow_rate', 'Q'),
                            ('channel_water_x-section__mean_depth', 'd'),
                            ('channel_water_x-section__volume_flux', 'u'),
                            ('channel_water__volume', 'vol'),
                            ('channel_water_flow__half_of_fanning_friction_factor', 'f'),
                            ('channel_water_x-section__hydraulic_radius', 'Rh'),
                            ('channel_water_surface__slope', 'S_free'),
                            ('channel_bottom_surface__slope', 'S_bed'),
                            ('channel_x-section_trapezoid_bottom__width', 'width')):
        assert field_err(comp.get_values(long_name), getattr(ch, attr)) <= RTOL, attr
    for attr in ('Q_outlet', 'u_outlet', 'd_outlet', 'f_outlet', 'Q_peak', 'T_peak', 'u_peak',
                 'd_peak', 'vol_Q', 'vol_R'):
        want = float(getattr(ch, attr))
        assert abs(float(getattr(comp, attr)) - want) <= RTOL * max(abs(want), 1e-30), attr
    comp.finalize()
    o = ch.mins_and_maxes()
    assert abs(float(comp.Q_max) - o['Q_max']) <= RTOL * o['Q_max']
    if wave == 'kinematic':
        # the reference's init log: 144 zero slopes replaced by the smallest positive one
        assert float('%.12g' % comp.S_bed.min()) == log['S_min']


def test_failure_path_sets_status_and_raises(golden_dir, tmp_path):
    from topoflow_old_b200.components import channels_kinematic_wave
    d = stage(golden_dir, 'treynor', tmp_path)
    comp = channels_kinematic_wave.channels_component()
    comp.initialize(cfg_file=os.path.join(d, 'June_20_67_channels_kinematic_wave.cfg'), SILENT=True)
    connect_disabled_providers(comp, och.scalar(np.nan))
    with pytest.raises(RuntimeError, match='Negative depth found.'):
        comp.update(-1.0)
    comp.finalize()


def test_d8_component_standalone_area(golden_dir, tmp_path):
    from topoflow_old_b200.components import d8_global
    d = stage(golden_dir, 'treynor', tmp_path)
    c = d8_global.d8_component()
    c.site_prefix, c.case_prefix, c.in_directory = 'Treynor', 'June_20_67', d + os.sep
    c.initialize(cfg_file=None, SILENT=True, REPORT=False)
    c.update(SILENT=True)
    want = np.fromfile(os.path.join(d, 'Treynor_area.rtg'), dtype='>f4').reshape(44, 29)
    assert np.array_equal(c.A.view(np.uint32), want.astype(np.float32).view(np.uint32))
    assert c.resolve.sum() == 10161
    c.finalize()


def test_update_until_and_unchanged_forcing_take_the_fast_path(golden_dir, tmp_path):
    """update_until() queues the steps between two output times as one device
    call, and a rain grid that has not changed is not uploaded again: same
    files, same scalars as 600 single update() calls."""
    from topoflow_old_b200.components import channels_kinematic_wave
    runs = {}
    for mode in ('stepwise', 'batched'):
        d = stage(golden_dir, 'plane_canal', tmp_path / mode)
        rain = np.fromfile(os.path.join(d, 'Rain_100mmph_on_right.rtg'), dtype='<f4').reshape(20, 20)
        P = np.float64(rain * np.float32(np.float64(1) / np.float64(3600000)))
        comp = channels_kinematic_wave.channels_component()
        comp.initialize(cfg_file=os.path.join(d, 'Case1_channels_kinematic_wave.cfg'), SILENT=True)
        connect_disabled_providers(comp, P)
        uploads = []
        real = comp.d8.strip.set_forcing
        comp.d8.strip.set_forcing = lambda k, v: (uploads.append(k), real(k, v))[1]
        if mode == 'stepwise':
            for _ in range(600):
                comp.update(-1.0)
        else:
            comp.update_until(3600.0)
        assert comp.get_status() == 'updated' and float(comp.get_current_time()) == 3600.0
        assert comp.time_index == 600
        # the six fluxes went to the device once, not 600 times
        assert sorted(uploads) == [0, 1, 2, 3, 4, 5]
        runs[mode] = dict(Q=comp.get_values('channel_water_x-section__volume_flow_rate').copy(),
                          scal=[float(getattr(comp, n)) for n in ('Q_outlet', 'Q_peak', 'T_peak', 'u_peak', 'd_peak',
                                                                 'vol_Q', 'vol_R')], d=d)
        comp.finalize()
    a, b = runs['stepwise'], runs['batched']
    assert np.array_equal(a['Q'], b['Q'])
    assert a['scal'][:6] == b['scal'][:6]
    assert abs(a['scal'][6] - b['scal'][6]) <= 1e-12 * abs(a['scal'][6])
    for f in ('Case1_0D-Q.txt', 'Case1_0D-d.txt', 'Case1_0D-u.txt', 'Case1_2D-Q.rts'):
        assert open(os.path.join(a['d'], f), 'rb').read() == open(os.path.join(b['d'], f), 'rb').read(), f
