import os, sys, shutil, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import replay
from tests.test_gpu_components import stage
from topoflow_old_b200.components import channels_kinematic_wave
from topoflow_old_b200 import _lib as L
import pathlib
golden = os.path.join(ROOT, 'tests', 'golden')
tmp = pathlib.Path(tempfile.mkdtemp())
d = stage(golden, 'plane_canal', tmp)
comp = channels_kinematic_wave.channels_component()
cfg = os.path.join(d, 'Case1_channels_kinematic_wave.cfg')
np.set_printoptions(linewidth=250, precision=4)
try:
    fw, chan = replay.run_plane(golden, lambda div: replay.BmiPlaneChannels(comp, cfg, div))
    print('ran fine')
except Exception as e:
    print('EXC', e, 'time', float(comp.time), 'index', comp.time_index)
    s = comp.d8.strip
    for n, f in (('d', L.F_D), ('vol', L.F_VOL), ('u', L.F_U), ('Qnext', L.F_QNEXT), ('Q', L.F_Q)):
        g = s.get(f)
        print(n, 'nan count', np.isnan(g).sum(), 'where', np.argwhere(np.isnan(g))[:10].tolist())
    try:
        g = s.get(L.F_R); print('R nan', np.isnan(g).sum(), 'R range', np.nanmin(g), np.nanmax(g))
    except Exception as e2:
        print('R:', e2)
    print('code\n', s.d8_get(L.D8_CODE))
