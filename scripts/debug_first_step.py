"""First step vs the oracle (vol, d must be bit-identical): where not?"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests.cases import make_case, oracle_for          # noqa: E402
from tests.test_gpu_channels import gpu_for            # noqa: E402
from topoflow_old_b200 import _lib as L                # noqa: E402

nx, ny = int(sys.argv[1]), int(sys.argv[2])
d0 = float(sys.argv[3]) if len(sys.argv) > 3 else 0.02
c = make_case(nx, ny, seed=3, d0=d0)
for mode in ('march', 'tile'):
    os.environ['TF_B200_MARCH'] = '0' if mode == 'tile' else '1'
    ch = oracle_for(c, 0, 2.0)
    s = gpu_for(c, 0, 2.0)
    v0 = s.get(L.F_VOL)
    print(mode, 'initial vol equal:', np.array_equal(v0, ch.vol), 'initial d equal:', np.array_equal(s.get(L.F_D), ch.d))
    s.set_R(float(c['P_rain']))
    s.step(1)
    ch.update()
    for n, f in (('vol', L.F_VOL), ('d', L.F_D), ('u', L.F_U)):
        a, b = s.get(f), getattr(ch, n)
        bad = a != b
        if bad.any():
            rr, cc = np.nonzero(bad)
            rel = np.abs(a[bad] - b[bad]) / np.maximum(np.abs(b[bad]), 1e-300)
            print(mode, n, bad.sum(), 'differ; rows', rr.min(), rr.max(), 'cols', cc.min(), cc.max(), 'max rel', rel.max(),
                  'first', (rr[0], cc[0]), repr(a[rr[0], cc[0]]), repr(b[rr[0], cc[0]]),
                  'code', c['d8']['d8_grid'][rr[0], cc[0]], 'angle', np.float64(c['angle'])[rr[0], cc[0]] if np.ndim(c['angle']) else c['angle'],
                  'width', np.float64(c['width'])[rr[0], cc[0]])
        else:
            print(mode, n, 'identical')
    s.close()
