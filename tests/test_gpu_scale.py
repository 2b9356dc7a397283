"""Routing parity at sizes where the device layout is stressed: rows that span
hundreds of column bands / tiles, byte offsets beyond 2^31, segment and band
seams of the marching kernels, the look-ahead of the tile kernels.

The NumPy oracle does about 1e7 cell-updates/s, so 2048^2 x 10 steps per wave
and 4096^2 x 3 steps are a few tens of seconds each.  Same tolerance as
tests/test_gpu_channels.py (1e-10 relative, north star); vol and d of the
first step are compared bit for bit.
"""
import os

import numpy as np
import pytest

from tests.cases import make_case, oracle_for
from tests.test_gpu_channels import RTOL, WAVES, compare, field_err, gpu_for

pytestmark = pytest.mark.gpu

_CASES = {}


def _case(nx, ny, **kw):
    key = (nx, ny, tuple(sorted(kw.items())))
    if key not in _CASES:
        _CASES.clear()                       # one big case in memory at a time
        _CASES[key] = make_case(nx, ny, **kw)
    return _CASES[key]


def _route(c, wave, n_steps, d0_check=True):
    from topoflow_old_b200 import _lib as L
    dt = 0.5 if wave == 'dynamic' else 2.0
    ch = oracle_for(c, WAVES[wave], dt)
    s = gpu_for(c, WAVES[wave], dt)
    try:
        s.set_R(float(c['P_rain']))
        taps, _, bad = s.step(1)
        ch.update()
        assert bad[0] == -1 and bad[1] == -1
        if d0_check:                         # nothing inexact has fed back yet
            assert np.array_equal(s.get(L.F_VOL), ch.vol)
            assert np.array_equal(s.get(L.F_D), ch.d)
        taps, _, bad = s.step(n_steps - 1)
        for _ in range(n_steps - 1):
            ch.update()
        assert bad[0] == -1 and bad[1] == -1
        # The dynamic wave forms u = M / vol with M a sum of momentum terms that
        # nearly cancel in a few cells of four million (u ~ 1e-6 of the field's
        # maximum, from terms 1e4 times larger: measured, scripts/debug_dyn.py).
        # There the 6e-14 relative difference that d has picked up by then (the
        # friction factor's cube root differs from glibc's pow in the last bits,
        # and pow has its own error) is amplified, through (d - d_parent)/ds,
        # to 1e-9 of that tiny velocity: 2 cells of 4.2 million exceed 1e-10 of
        # their own value, by an absolute 2e-15 m/s.  Those cells are held to
        # 1e-14 of the field maximum (floor 1e-4) instead; vol, d and Q keep the
        # usual bound everywhere, and u keeps it wherever |u| > 1e-4 max|u|.
        worst = compare(s, ch, '%s %dx%d' % (wave, c['nx'], c['ny']),
                        u_floor=1e-4 if wave == 'dynamic' else 1e-9)
        assert field_err(taps[-1, 0], np.array([ch.Q_outlet, ch.u_outlet, ch.d_outlet, ch.f_outlet],
                                               dtype=np.float64).reshape(4)) <= RTOL
        assert ch.Q.max() > 0.0
        return worst
    finally:
        s.close()


@pytest.mark.parametrize('wave', ['kinematic', 'diffusive', 'dynamic'])
def test_2048_square_ten_steps(wave):
    c = _case(2048, 2048, seed=3, d0=0.02)
    _route(c, wave, 10)


def test_4096_square_kinematic_three_steps():
    c = _case(4096, 4096, seed=4, d0=0.02)
    _route(c, 'kinematic', 3)


@pytest.mark.parametrize('wave', ['kinematic', 'dynamic'])
def test_ragged_wide_grid(wave):
    """nx = 16384 + 17 (not a multiple of anything), 64 rows: 512+ tiles /
    274 bands per row, a last band that overhangs the row end."""
    c = _case(16384 + 17, 64, seed=5, d0=0.02)
    _route(c, wave, 12)


@pytest.mark.parametrize('wave', ['kinematic', 'diffusive', 'dynamic'])
def test_marching_and_tile_kernels_agree_bit_for_bit(wave):
    """TF_B200_MARCH=1 runs the packed layout on the marching kernels, which
    fuse the two phases of the diffusive / dynamic step; the default is the
    tile kernels.  Both evaluate the same expressions in the same order, so
    every field must agree exactly, seams of bands (60 / 64 columns) and of
    segments (TF_B200_SEG_ROWS rows) included."""
    from topoflow_old_b200 import _lib as L
    c = _case(333, 211, seed=9, d0=0.01)
    dt = 0.5 if wave == 'dynamic' else 2.0
    out = {}
    for mode in ('march', 'march_seg7', 'tile'):
        os.environ.pop('TF_B200_MARCH', None)
        os.environ.pop('TF_B200_SEG_ROWS', None)
        if mode != 'tile':
            os.environ['TF_B200_MARCH'] = '1'
        if mode == 'march_seg7':
            os.environ['TF_B200_SEG_ROWS'] = '7'
        try:
            s = gpu_for(c, WAVES[wave], dt)
        finally:
            os.environ.pop('TF_B200_MARCH', None)
            os.environ.pop('TF_B200_SEG_ROWS', None)
        try:
            s.set_R(float(c['P_rain']))
            taps, _, bad = s.step(150)
            assert bad[0] == -1 and bad[1] == -1
            out[mode] = {n: s.get(f) for n, f in (('vol', L.F_VOL), ('d', L.F_D), ('u', L.F_U),
                                                    ('Q', L.F_Q), ('mut', L.F_MUTATED))}
            out[mode]['taps'] = taps
        finally:
            s.close()
    for mode in ('march', 'march_seg7'):
        for k, v in out['tile'].items():
            assert np.array_equal(out[mode][k], v), (mode, k)
    assert out['tile']['Q'].max() > 0


def test_second_init_on_one_handle_equals_a_fresh_handle():
    """tf_chan_init on a handle that has already run: u, Q, R and the overbank
    state restart as the reference's initialize() does (channels_base.py:
    900-910), whatever the first run left behind."""
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200.routing import setup_channels
    c = _case(96, 80, seed=11)
    cp = lambda v: v.copy() if np.ndim(v) else float(v)

    def setup(s, wave, dt):
        setup_channels(s, wave, c['law'], dt, c['da'], c['slope'].copy(), cp(c['width']),
                       cp(c['angle']), cp(c['rough']), outlets=c['outlets'])

    s = gpu_for(c, L.DYNAMIC, 0.5)
    fresh = gpu_for(c, L.DIFFUSIVE, 2.0)
    try:
        s.set_R(np.full((c['ny'], c['nx']), 2e-5))          # a grid R and 40 dynamic steps first
        s.step(40)
        setup(s, L.DIFFUSIVE, 2.0)                           # ... then a second initialisation
        for h in (s, fresh):
            h.set_R(float(c['P_rain']))
            _, _, bad = h.step(60)
            assert bad[0] == -1 and bad[1] == -1
        for f in (L.F_VOL, L.F_D, L.F_U, L.F_Q, L.F_MUTATED):
            assert np.array_equal(s.get(f), fresh.get(f)), f
    finally:
        s.close()
        fresh.close()
