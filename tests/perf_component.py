#!/usr/bin/env python
"""Per-update cost of the BMI component on the reference's own small case
(Treynor, 44 x 29, configs[0] of BASELINE.json), next to the NumPy oracle on
the same host: what a user switching the component inside EMELI would see."""
import os
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))   # lives under tests/: it times the oracle next to the component
sys.path.insert(0, ROOT)
from oracle import channels as och          # noqa: E402  (baseline leg only)
from tests import replay                    # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
    gd = os.path.join(ROOT, 'tests', 'golden')
    out = {}
    for wave in ('kinematic', 'diffusive', 'dynamic'):
        import importlib
        mod = importlib.import_module('topoflow_old_b200.components.channels_%s_wave' % wave)
        d = os.path.join(tempfile.mkdtemp(), 'treynor')
        shutil.copytree(os.path.join(gd, 'treynor'), d)
        for f in os.listdir(d):
            if f.endswith('_path_info.cfg'):
                p = os.path.join(d, f)
                txt = open(p).read().replace('~/TF_Output/Treynor', '.')
                open(p, 'w').write(txt)
        comp = mod.channels_component()
        stdout = sys.stdout
        sys.stdout = open(os.devnull, 'w')
        comp.initialize(cfg_file=os.path.join(d, 'June_20_67_channels_%s_wave.cfg' % wave), SILENT=True)
        sys.stdout = stdout
        P = och.scalar(60.0 / 3.6e6)
        comp.set_values('atmosphere_water__rainfall_volume_flux', P)
        for name in ('snowpack__melt_volume_flux', 'land_surface_water__baseflow_volume_flux',
                     'glacier_ice__melt_volume_flux', 'land_surface_water__evaporation_volume_flux',
                     'soil_surface_water__infiltration_volume_flux'):
            comp.set_values(name, och.scalar(0))
        for name in ('canals__count', 'sinks__count', 'sources__count'):
            comp.set_values(name, np.array(0, dtype='int32'))
        comp.SAVE_Q_GRIDS = comp.SAVE_U_GRIDS = comp.SAVE_D_GRIDS = comp.SAVE_F_GRIDS = False
        for _ in range(100):
            comp.update(-1.0)
        t = time.perf_counter()
        for _ in range(n):
            comp.update(-1.0)
        gpu_us = (time.perf_counter() - t) / n * 1e6
        ch = replay.make_oracle(gd, wave)
        ch.P_rain = P
        for _ in range(50):
            ch.update()
        t = time.perf_counter()
        for _ in range(n // 4):
            ch.update()
        cpu_us = (time.perf_counter() - t) / (n // 4) * 1e6
        out[wave] = (gpu_us, cpu_us)
        print('%-10s component.update() %8.1f us   NumPy oracle update() %8.1f us   ratio %.2f'
              % (wave, gpu_us, cpu_us, cpu_us / gpu_us), flush=True)
        sys.stdout = open(os.devnull, 'w')
        comp.finalize()
        sys.stdout = stdout


def big(n):
    """The plane case blown up to n x n cells (same cfg files, a fractal DEM and
    its D8 slopes written as RTG): per-update cost of the component with a 0-d
    rain rate, with a constant rain GRID handed over before every update (what
    EMELI does), and with update_until batching."""
    from topoflow_old_b200.components import channels_kinematic_wave
    from topoflow_old_b200.utils import synth
    from oracle import d8 as od8      # noqa: F401  (only to keep the import surface of this script in one place)
    gd = os.path.join(ROOT, 'tests', 'golden')
    d = os.path.join(tempfile.mkdtemp(), 'plane_canal')
    shutil.copytree(os.path.join(gd, 'plane_canal'), d)
    rti = os.path.join(d, 'plane.rti')
    txt = open(rti).read().replace('Number of columns:  20', 'Number of columns:  %d' % n)
    txt = txt.replace('Number of rows:     20', 'Number of rows:     %d' % n)
    open(rti, 'w').write(txt)
    dem = synth.fractal_dem(n, n, seed=3, cell=10.0, roughness=8.0)
    dem.astype('<f4').tofile(os.path.join(d, 'plane_DEM.rtg'))
    np.full((n, n), 0.01, dtype='<f4').tofile(os.path.join(d, 'plane_slope.rtg'))
    res = {}
    for mode in ('scalar rain', 'constant rain grid', 'constant rain grid, update_until'):
        comp = channels_kinematic_wave.channels_component()
        stdout = sys.stdout
        sys.stdout = open(os.devnull, 'w')
        comp.initialize(cfg_file=os.path.join(d, 'Case1_channels_kinematic_wave.cfg'), SILENT=True)
        sys.stdout = stdout
        P = och.scalar(60.0 / 3.6e6) if mode == 'scalar rain' else np.full((n, n), 60.0 / 3.6e6)
        for name in ('snowpack__melt_volume_flux', 'land_surface_water__baseflow_volume_flux',
                     'glacier_ice__melt_volume_flux', 'land_surface_water__evaporation_volume_flux',
                     'soil_surface_water__infiltration_volume_flux'):
            comp.set_values(name, och.scalar(0))
        for name in ('canals__count', 'sinks__count', 'sources__count'):
            comp.set_values(name, np.array(0, dtype='int32'))
        for k in 'QUDF':
            setattr(comp, 'SAVE_%s_GRIDS' % k, False)
            setattr(comp, 'SAVE_%s_PIXELS' % k, False)
        steps = 200
        comp.set_values('atmosphere_water__rainfall_volume_flux', P)
        for _ in range(20):
            comp.update(-1.0)
        t = time.perf_counter()
        if mode.endswith('update_until'):
            comp.set_values('atmosphere_water__rainfall_volume_flux', P)
            comp.update_until(float(comp.time) + steps * float(comp.dt))
        else:
            for _ in range(steps):
                comp.set_values('atmosphere_water__rainfall_volume_flux', P)     # emeli.py:1841-1933
                comp.update(-1.0)
        res[mode] = (time.perf_counter() - t) / steps * 1e6
        print('%dx%d  %-34s component.update() %9.1f us/step' % (n, n, mode, res[mode]), flush=True)
        sys.stdout = open(os.devnull, 'w')
        comp.finalize()
        sys.stdout = stdout
    print('constant rain grid / scalar rain = %.3f' % (res['constant rain grid'] / res['scalar rain']))


if __name__ == '__main__':
    if len(sys.argv) > 2 and sys.argv[1] == 'big':
        big(int(sys.argv[2]))
    else:
        main()
