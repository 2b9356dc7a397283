"""CPU oracle for the TopoFlow 3.5 D8 + channel-routing hot path.

TEST INFRASTRUCTURE ONLY.  This package is a NumPy restatement of the
reference's algorithm (peckhams/topoflow_old, topoflow/components/
channels_*.py, d8_base.py, d8_global.py).  It exists so the CUDA path can be
checked against it.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it; the
product package ``topoflow_old_b200`` never does.

Parity status: PINNED.
  * D8 codes      : bit-exact vs the reference's shipped Treynor_flow.rtg
  * D8 area       : bit-exact (float32) vs Treynor_area.rtg
  * tie-break LUT : equals the RT3 16x16 table, sum 10161
                    (components/tests/test_d8_base_UNFINISHED.py:101-125)
  * routing       : reproduces the shipped Test_Plane_Canal run in full -- rain
                    grid + the 240 s canal, replayed through oracle/coupling.py:
                    every column of Case1_0D-{Q,d,u}.txt digit for digit and
                    all 60 frames of Case1_2D-Q.rts bit for bit (this pins
                    update_diversions, channels_base.py:1226-1297) -- and, driven through a restatement of the EMELI loop
                    (oracle/coupling.py), the shipped Treynor runs Test1_kin_wave,
                    Test2_diff_wave and Test2_diff_wave_dt=6s: hydrographs digit
                    for digit, every frame of their float32 Q grid stacks
                    (158-159 frames x 1276 cells) bit for bit, Q/u/d peaks and the
                    final report (Q_final, vol_Q, vol_R at the driver's stop)
                    equal to the console log to its 12 printed digits;
                    Test8_snow_degree_day (snowmelt as a second source, through
                    EMELI's interpolation of a rebound scalar) likewise, with
                    vol_SM; Test12_dynam_wave identical through the peak, within 1e-5
                    on the recession (glibc vs the 2017 macOS libm)
see tests/test_oracle_*.py.
"""
