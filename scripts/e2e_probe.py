"""Where does the per-call overhead of tf_chan_step go?  (diagnostic)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
class A: pass
a = A(); a.size = int(sys.argv[1]) if len(sys.argv) > 1 else 8192; a.wave = 'kinematic'; a.scaling = 'weak'; a.generic = False; a.fill_pits = False
s = bench.build_strip(a, 0, 1, 0, quiet=True)
s.step_async(5); s.sync(); s.timing()
def t(f, n):
    s.sync(); t0 = time.perf_counter(); f(); s.sync(); return (time.perf_counter() - t0) / n * 1e3
print('step_async(20)+sync      %.4f ms/step' % t(lambda: s.step_async(20), 20))
print('step(20) (taps recorded) %.4f ms/step' % t(lambda: s.step(20), 20))
print('20 x step(1)             %.4f ms/step' % t(lambda: [s.step(1) for _ in range(20)], 20))
print('20 x (step_async(1)+sync)%.4f ms/step' % t(lambda: [(s.step_async(1), s.sync()) for _ in range(20)], 20))
ms, n, _, _ = s.timing()
