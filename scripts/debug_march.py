"""Marching kernels vs tile kernels on one small case: where do they differ?
(diagnostic for the GPU box; prints a summary per wave and field)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests.cases import make_case                      # noqa: E402
from tests.test_gpu_channels import gpu_for            # noqa: E402
from topoflow_old_b200 import _lib as L                # noqa: E402

nx, ny = int(sys.argv[1]) if len(sys.argv) > 1 else 333, int(sys.argv[2]) if len(sys.argv) > 2 else 211
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
c = make_case(nx, ny, seed=9, d0=0.01)
for wave, dt in ((0, 2.0), (1, 2.0), (2, 0.5)):
    out = {}
    for mode in ('march', 'tile'):
        os.environ['TF_B200_MARCH'] = '0' if mode == 'tile' else '1'
        s = gpu_for(c, wave, dt)
        s.set_R(float(c['P_rain']))
        res = []
        for k in range(steps):
            _, _, bad = s.step(1)
            res.append({n: s.get(f) for n, f in (('vol', L.F_VOL), ('d', L.F_D), ('u', L.F_U), ('Q', L.F_QNEXT))})
        out[mode] = res
        s.close()
    for k in range(steps):
        for n in ('vol', 'd', 'u', 'Q'):
            a, b = out['march'][k][n], out['tile'][k][n]
            bad = ~((a == b) | (np.isnan(a) & np.isnan(b)))
            if bad.any():
                rr, cc = np.nonzero(bad)
                print('wave %d step %d %s: %d cells differ; rows %d..%d cols %d..%d; first (%d,%d) march %r tile %r; '
                      'cols mod 60: %s' % (wave, k + 1, n, bad.sum(), rr.min(), rr.max(), cc.min(), cc.max(),
                                           rr[0], cc[0], a[rr[0], cc[0]], b[rr[0], cc[0]],
                                           np.unique(cc % 60)[:12]))
            else:
                print('wave %d step %d %s: identical' % (wave, k + 1, n))
