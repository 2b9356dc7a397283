"""Pin the routing oracle against the reference's shipped Treynor hydrographs:
the same DEM, channel grids, rain series and Green-Ampt parameters, driven
through a restatement of EMELI's loop and time interpolation, must print the
numbers in 00_Output_Files/Test*/June_20_67_0D-Q.txt and the peak values of
00_Full_Output.txt."""
import json
import os

import numpy as np
import pytest

from tests import replay


@pytest.mark.parametrize('wave', ['kinematic', 'diffusive', 'diffusive_dt6', 'dynamic'])
def test_treynor_hydrograph_matches_reference_output(golden_dir, wave):
    gold = replay.golden_rows(golden_dir, wave)
    log = json.load(open(os.path.join(golden_dir, 'treynor', 'log_scalars.json')))[replay.CASES[wave][1]]
    ch = replay.make_oracle(golden_dir, wave)
    chan = replay.OracleChannels(ch)
    fw = replay.run(golden_dir, chan, replay.CASES[wave][3])
    rows = np.array(chan.rows)[:len(gold)]
    assert rows.shape == gold.shape and gold.shape[0] >= 158
    assert np.array_equal(rows[:, 0], gold[:, 0])                      # time column
    if wave != 'dynamic':
        assert np.array_equal(rows, gold)                              # every printed digit
        # ... and every cell of every frame of the float32 grid stack, bit for bit
        stack = replay.golden_stack(golden_dir, wave)
        mine = np.array(chan.frames[:len(stack)])
        assert stack.shape[0] >= 158 and mine.shape == stack.shape
        assert np.array_equal(mine.view(np.uint32), stack.view(np.uint32))
    else:
        # dynamic wave, dt = 0.5 s, 16 800 steps: identical through the rising limb and the
        # peak; on the recession the 2017 run (macOS libm pow) and glibc drift apart by ~1e-6
        early = gold[:, 0] <= 90.0
        assert np.array_equal(rows[early], gold[early])
        assert np.max(np.abs(rows - gold)) < 1e-5
    # peaks printed in 00_Full_Output.txt (12 significant digits)
    for name, attr in (('Q_peak', 'Q_peak'), ('u_peak', 'u_peak'), ('d_peak', 'd_peak'), ('T_peak', 'T_peak')):
        assert float('%.12g' % float(getattr(ch, attr))) == log[name], name
    # the final report of the run (topoflow_driver.print_final_report): discharge at the stop,
    # integrated discharge (update_Q_out_integral) and integrated runoff (update_R_integral)
    if wave != 'dynamic':
        replay.run_to_the_logged_stop(fw, chan, wave, log)
        for name, attr in (('Q_final', 'Q_outlet'), ('vol_Q', 'vol_Q'), ('vol_R', 'vol_R')):
            assert float('%.12g' % float(getattr(ch, attr))) == log[name], name


def test_treynor_with_degree_day_snow_matches_reference_output(golden_dir):
    """Test8_snow_degree_day: the same run with snowmelt as a second source of
    runoff.  SM reaches the channels through EMELI's linear interpolation of a
    REBOUND 0-d array (a ramp over the snow component's first hour), so this
    pins the coupling restatement as well as the provider."""
    case = 'kinematic_snow'
    gold = replay.golden_rows(golden_dir, case)
    log = json.load(open(os.path.join(golden_dir, 'treynor', 'log_scalars.json')))[replay.CASES[case][1]]
    ch = replay.make_oracle(golden_dir, case)
    chan = replay.OracleChannels(ch)
    snow = replay.degree_day_snow(golden_dir)
    fw = replay.run(golden_dir, chan, replay.CASES[case][3], snow=snow)
    rows = np.array(chan.rows)[:len(gold)]
    assert gold.shape == (159, 6) and np.array_equal(rows, gold)
    assert gold[1, 1] > 0                      # melt water reaches the outlet before any rain does
    for name in ('Q_peak', 'u_peak', 'd_peak', 'T_peak'):
        assert float('%.12g' % float(getattr(ch, name))) == log[name], name
    replay.run_to_the_logged_stop(fw, chan, case, log)
    for name, attr in (('Q_final', 'Q_outlet'), ('vol_Q', 'vol_Q'), ('vol_R', 'vol_R')):
        assert float('%.12g' % float(getattr(ch, attr))) == log[name], name
    assert float('%.10g' % float(snow.vol_SM)) == log['vol_SM']
