import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests.cases import make_case, oracle_for
from tests.test_gpu_channels import gpu_for
from topoflow_old_b200 import _lib as L
n = int(sys.argv[1])
c = make_case(n, n, seed=3, d0=0.02)
ch = oracle_for(c, 2, 0.5)
for mode in ('march', 'tile'):
    os.environ['TF_B200_MARCH'] = '0' if mode == 'tile' else '1'
    s = gpu_for(c, 2, 0.5)
    s.set_R(float(c['P_rain']))
    s.step(10)
    if mode == 'march':
        for _ in range(10):
            ch.update()
    u, ur = s.get(L.F_U), ch.u
    scale = np.abs(ur).max()
    err = np.abs(u - ur) / np.maximum(np.abs(ur), 1e-9 * scale)
    idx = np.argsort(err.ravel())[::-1][:8]
    print(mode, 'scale', scale, 'n cells > 1e-10:', int((err > 1e-10).sum()))
    for i in idx:
        r, cc = divmod(i, n)
        print('  cell', (r, cc), 'err', err[r, cc], 'got', repr(u[r, cc]), 'ref', repr(ur[r, cc]), 'd', ch.d[r, cc], 'vol', ch.vol[r, cc], 'code', c['d8']['d8_grid'][r, cc])
    s.close()
