import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests.cases import make_case, oracle_for
from tests.test_gpu_channels import gpu_for
from topoflow_old_b200 import _lib as L
nx, ny = int(sys.argv[1]), int(sys.argv[2])
c = make_case(nx, ny, seed=3, d0=0.0)
ch = oracle_for(c, 0, 2.0)
ch.update()
for rep in range(2):
    s = gpu_for(c, 0, 2.0)
    s.set_R(float(c['P_rain']))
    s.step(1)
    a, b = s.get(L.F_D), ch.d
    bad = a != b
    rr, cc = np.nonzero(bad)
    print(nx, ny, 'rep', rep, 'bad', bad.sum())
    if bad.any():
        print(' bands', np.unique(cc // 64)[:40])
        print(' col%64', np.unique(cc % 64)[:64])
        print(' row%64 hist', np.bincount(rr % 64, minlength=64))
        print(' rows/64', np.unique(rr // 64)[:40])
        print(' gpu values uniq', np.unique(a[bad])[:5], 'code of bad', np.unique(c['d8']['d8_grid'][bad]))
    s.close()
