#!/usr/bin/env python
"""Time the D8 builder and the contributing-area kernel (reported
separately from the routing metric, as cells/s and seconds)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from topoflow_old_b200 import _lib as L          # noqa: E402
from topoflow_old_b200.engine import Strip       # noqa: E402
from topoflow_old_b200.utils import synth        # noqa: E402


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [4096]
    for n in sizes:
        from concurrent.futures import ThreadPoolExecutor
        chunks = [(a, min(a + 512, n)) for a in range(0, n, 512)]
        with ThreadPoolExecutor(16) as ex:
            dem = np.concatenate(list(ex.map(
                lambda ab: synth.fractal_dem(n, n, row0=ab[0], nrows=ab[1] - ab[0], roughness=8.0), chunks)))
        ones = np.ones(n, dtype=np.float32)
        with Strip(n, n) as s:
            t0 = time.perf_counter()
            stats = s.d8_build(dem, 30 * ones, 30 * ones, np.float32(np.sqrt(1800.0)) * ones)
            t_build = time.perf_counter() - t0
            t0 = time.perf_counter()
            stats2 = s.d8_build(dem, 30 * ones, 30 * ones, np.float32(np.sqrt(1800.0)) * ones)
            t_build2 = time.perf_counter() - t0
            t0 = time.perf_counter()
            n_bad = s.d8_area(900.0, units_km=True)
            t_area = time.perf_counter() - t0
            A = s.d8_get(L.D8_AREA)
            print(json.dumps({'size': n, 'd8_build_s_first': t_build, 'd8_build_s': t_build2,
                              'd8_build_cells_per_s': n * n / t_build2, 'stats': stats2,
                              'area_s': t_area, 'area_cells_per_s': n * n / t_area,
                              'area_unresolved': n_bad, 'A_max_km2': float(A.max()),
                              'note': 'host->device copy of the DEM is inside d8_build'}), flush=True)


if __name__ == '__main__':
    main()
