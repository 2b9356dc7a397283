"""The CUDA channels component inside the replayed Treynor runs: the text
hydrograph it writes must equal the file the reference shipped (7 printed
decimals), and its peak values the numbers in the reference's console log."""
import importlib
import json
import os
import re

import numpy as np
import pytest

from tests import replay
from tests.test_gpu_components import stage

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('case_key', ['kinematic', 'diffusive', 'diffusive_dt6', 'dynamic'])
def test_component_reproduces_shipped_hydrograph(golden_dir, tmp_path, case_key):
    wave = replay.wave_of(case_key)
    mod = importlib.import_module('topoflow_old_b200.components.channels_%s_wave' % wave)
    d = stage(golden_dir, 'treynor', tmp_path)
    gold_dir, case, dt, minutes = replay.CASES[case_key]
    cfg = os.path.join(d, 'June_20_67_channels_%s_wave.cfg' % wave)
    txt = re.sub(r'^dt(\s*)\|[^|]*\|', 'dt\\1| %s |' % dt, open(cfg).read(), flags=re.M)
    open(cfg, 'w').write(txt)
    comp = mod.channels_component()
    comp.initialize(cfg_file=cfg, SILENT=True)
    assert float(comp.dt) == dt
    for name in ('land_surface_water__baseflow_volume_flux', 'glacier_ice__melt_volume_flux',
                 'land_surface_water__evaporation_volume_flux'):
        comp.set_values(name, np.array(0.0))
    comp.set_values('water-liquid__mass-per-volume_density', np.array(1000.0))
    for name in ('canals__count', 'sinks__count', 'sources__count'):
        comp.set_values(name, np.array(0, dtype='int32'))
    chan = replay.BmiChannels(comp)
    fw = replay.run(golden_dir, chan, minutes)
    log = json.load(open(os.path.join(golden_dir, 'treynor', 'log_scalars.json')))[case]
    for name in ('Q_peak', 'u_peak', 'd_peak'):
        got = float(getattr(comp, name))
        assert abs(got - log[name]) <= 1e-10 * log[name], (name, got, log[name])
    assert float('%.12g' % float(comp.T_peak)) == log['T_peak']
    if wave != 'dynamic':                       # the final report of the reference's run
        replay.run_to_the_logged_stop(fw, chan, case_key, log)
        for name, attr in (('Q_final', 'Q_outlet'), ('vol_Q', 'vol_Q'), ('vol_R', 'vol_R')):
            got = float(getattr(comp, attr))
            assert abs(got - log[name]) <= 1e-10 * log[name], (name, got, log[name])
    comp.finalize()
    mine = open(os.path.join(d, 'June_20_67_0D-Q.txt')).read().splitlines()
    gold = open(os.path.join(d, gold_dir, 'June_20_67_0D-Q.txt')).read().splitlines()
    assert mine[:2] == gold[:2]                                   # header and rule
    a = np.array([r.split() for r in mine[2:2 + len(gold) - 2]], dtype=float)
    b = np.array([r.split() for r in gold[2:]], dtype=float)
    assert a.shape == b.shape and np.array_equal(a[:, 0], b[:, 0])
    tol = 1.0000001e-7 if wave != 'dynamic' else 1e-5            # one unit of the last printed digit
    assert np.max(np.abs(a - b)) <= tol
    if wave != 'dynamic':
        assert np.mean(a == b) > 0.99                             # and almost always the same digit
        # the grid stack the component wrote (device-side float32 frames, MSB as the RTI says)
        gold_stack = replay.golden_stack(golden_dir, case_key)
        mine_stack = np.fromfile(os.path.join(d, 'June_20_67_2D-Q.rts'), dtype='>f4')
        mine_stack = mine_stack[:gold_stack.size].reshape(gold_stack.shape).astype(np.float32)
        ulp = np.abs(mine_stack.view(np.int32).astype(np.int64) - gold_stack.view(np.int32).astype(np.int64))
        assert ulp.max() <= 1 and np.mean(ulp == 0) > 0.99
