"""Pin the routing oracle against the reference's shipped hydrographs.

examples/Test_Plane_Canal/00_Output_Files/Case1_0D-{Q,d,u}.txt were written by
the reference itself (TopoFlow 3.5, kinematic wave, dt = 6 s, 60 min).  The
monitored cells in column 18 are neither upstream nor downstream of the canal,
so the channels component alone (constant rain grid from the met component,
everything else Disabled) must reproduce those columns digit for digit.
"""
import os

import numpy as np
import pytest

from oracle import channels as och
from oracle import d8 as od8

PTS = [(10, 18), (1, 18), (10, 16), (1, 16), (10, 3), (9, 3), (8, 3), (1, 3),
       (3, 5), (2, 5), (1, 5), (1, 1)]          # Case1_outlets.txt as (row, col)


def plane_inputs(golden_dir):
    g = os.path.join(golden_dir, 'plane_canal')
    rd = lambda f: np.fromfile(os.path.join(g, f), dtype='<f4').reshape(20, 20)
    dem, slope, rain = rd('plane_DEM.rtg'), rd('plane_slope.rtg'), rd('Rain_100mmph_on_right.rtg')
    # met_base.py:1799-1800 -> model_input.read_next(..., factor=mmph_to_mps):
    # float32 data scaled in place (float32 arithmetic under NumPy-1 casting),
    # then widened (model_input.py:158-172)
    P = np.float64(rain * np.float32(np.float64(1) / np.float64(3600000)))
    return dem, slope, P


def golden(golden_dir, var):
    path = os.path.join(golden_dir, 'plane_canal', 'out', 'Case1_0D-%s.txt' % var)
    rows = [l.split() for l in open(path).read().splitlines()[2:]]
    return np.array(rows, dtype=float)


def run_plane(golden_dir, n_steps=600):
    dem, slope, P = plane_inputs(golden_dir)
    b = od8.build(dem, 10.0, 10.0)
    ch = och.ChannelsOracle(b, och.KINEMATIC, dt=6.0, da=np.float64(10.0) * np.float64(10.0),
                            slope=np.float64(slope), width=10.0, angle=0.0,
                            nval=0.3, outlet_ID=PTS[0])
    ch.P_rain = P
    rows = {v: [] for v in 'Qdu'}
    for _ in range(n_steps):
        t_write = float(ch.time)            # write_output_files runs before update_time
        ch.update()
        if int(t_write) % 30 == 0:          # channels_base.py:2481-2491
            for v in 'Qdu':
                # text_ts_files.py:149-186: float32 values, '%15.7f'
                vals = np.float32(getattr(ch, v)[tuple(np.array(PTS).T)])
                rows[v].append([float('%15.7f' % (t_write / 60.0))] +
                               [float('%15.7f' % x) for x in vals])
    return ch, {v: np.array(r) for v, r in rows.items()}


def test_plane_hydrographs_match_reference_output(golden_dir):
    """The routing alone (no canal): the columns the canal cannot reach."""
    ch, rows = run_plane(golden_dir)
    for v in 'Qdu':
        g = golden(golden_dir, v)
        assert g.shape == (120, 13) and rows[v].shape == g.shape
        assert np.array_equal(rows[v][:, 0], g[:, 0])          # time column
        for col in (1, 2):                                     # cells (10,18), (1,18)
            assert np.array_equal(rows[v][:, col], g[:, col]), (v, col)
    # analytic steady states in 00_README_Test_Plan_Canal.txt (F5-F7)
    assert abs(rows['Q'][-1, 1] - 0.025) < 1e-6
    assert abs(rows['Q'][-1, 2] - 0.05) < 2e-6


def test_plane_canal_run_replayed_in_full(golden_dir):
    """The shipped Case1 run as it was made: rain grid + the 240 s canal from
    (10,16) to (10,3) + EMELI's loop.  Every column of the three text files
    digit for digit, and all 60 frames of the float32 Q stack bit for bit."""
    from tests import replay

    class Recording(replay.OraclePlaneChannels):
        frames = []

        def update(self):
            t = float(self.ch.time)
            replay.OraclePlaneChannels.update(self)
            if int(t) % 60 == 0:                               # save_grid_dt = 60
                self.frames.append(np.float32(self.ch.Q))

    fw, chan = replay.run_plane(golden_dir, lambda div: Recording(golden_dir, div))
    for v in 'Qdu':
        g = golden(golden_dir, v)
        assert np.array_equal(np.array(chan.rows[v])[:120], g), v
    assert golden(golden_dir, 'Q')[:, 5].max() > 0.02          # the canal exit did carry water
    stack = np.fromfile(os.path.join(golden_dir, 'plane_canal', 'out', 'Case1_2D-Q.rts'),
                        dtype='<f4').reshape(60, 20, 20)
    mine = np.array(chan.frames[:60])
    assert np.array_equal(mine.view(np.uint32), stack.view(np.uint32))
    # what went in came out 40 updates later
    assert chan.ch.canals_out_Q[0] > 0 and fw.comps['div'].canal_Q.shape == (1, 40)


def test_mass_balance_and_checks(golden_dir):
    ch, _ = run_plane(golden_dir, n_steps=50)
    # vol_R integrates R*da*dt over the grid (channels_base.py:1151-1160)
    _, _, P = plane_inputs(golden_dir)
    assert np.isclose(float(ch.vol_R), 50 * np.sum(P * 100.0 * 6.0), rtol=1e-13)
    # all water is either still in channels or has piled up in border cells
    assert np.isclose(ch.vol.sum(), float(ch.vol_R), rtol=1e-12)
    ch.d[3, 3] = -1.0
    with pytest.raises(RuntimeError, match='Negative depth found.'):
        ch.check()
    ch.d[3, 3] = 0.0
    ch.u[4, 4] = np.nan
    with pytest.raises(RuntimeError, match='Negative or NaN velocity found.'):
        ch.check()


def treynor_inputs(golden_dir):
    g = os.path.join(golden_dir, 'treynor')
    rd = lambda f: np.fromfile(os.path.join(g, f), dtype='>f4').reshape(44, 29).astype(np.float32)
    dem = rd('Treynor_DEM.rtg')
    b = od8.build(dem, 30.0, 30.0)
    inp = och.prepare_inputs(b, rd('Treynor_slope.rtg'), rd('Treynor_chan-w.rtg'),
                             rd('Treynor_chan-a.rtg'), nval=rd('Treynor_chan-n.rtg'),
                             z0val=rd('Treynor_chan-z0.rtg'))
    return b, inp


def test_treynor_init_matches_log(golden_dir):
    """Counters the reference printed while initialising the kinematic case
    (Test1_kin_wave/00_Full_Output.txt)."""
    import json
    log = json.load(open(os.path.join(golden_dir, 'treynor', 'log_scalars.json')))['Test1_kin_wave']
    b, inp = treynor_inputs(golden_dir)
    ch = och.ChannelsOracle(b, och.KINEMATIC, dt=2.0, da=np.float64(900.0),
                            slope=inp['slope'], width=inp['width'], angle=inp['angle'],
                            nval=inp['nval'], outlet_ID=(34, 16))
    assert ch.n_zero_slope == log['n_zero_slope'] == 144
    assert float('%.12g' % ch.S_min) == log['S_min']
    assert float('%.12g' % ch.S_max) == log['S_max']
    assert float('%.12g' % ch.width.min()) == 2.0239880085      # "min(width)" in the log
    assert float('%.12g' % ch.width.max()) == 42.4264068604


@pytest.mark.parametrize('wave', [och.KINEMATIC, och.DIFFUSIVE, och.DYNAMIC])
@pytest.mark.parametrize('law', [False, True])
def test_treynor_all_variants_run_and_conserve(golden_dir, wave, law):
    b, inp = treynor_inputs(golden_dir)
    ch = och.ChannelsOracle(b, wave, dt=2.0 if wave != och.DYNAMIC else 0.5,
                            da=np.float64(900.0), slope=inp['slope'],
                            width=inp['width'].copy(), angle=inp['angle'].copy(),
                            nval=None if law else inp['nval'],
                            z0val=inp['z0val'] if law else None,
                            manning=not law, law_of_wall=law, outlet_ID=(34, 16))
    ch.P_rain = och.scalar(50.0 / 3.6e6)
    for _ in range(200):
        ch.update()
    assert ch.Q.max() > 0 and np.isfinite(ch.u).all()
    assert np.isclose(ch.vol.sum(), float(ch.vol_R), rtol=1e-10)
    assert not ch.u[b['noflow_IDs']].any() and not ch.d[b['noflow_IDs']].any()
