"""CUDA channel routing vs the oracle, through the C ABI.

Tolerance (BASELINE.json north_star): Q, d, u and the outlet hydrograph within
1e-10 relative in fp64.  `vol` and `d` use only correctly rounded operations
and start bit-identical; `u` goes through a cube root / log that differ from
glibc's pow / log by <= ~2 ulp, which then feeds back into vol.  The relative
error is taken per cell against max(|reference|, 1e-9 * max|reference field|):
cells that have drained to ~0 by cancellation carry no meaningful relative
error of their own.
"""
import numpy as np
import pytest

from oracle import channels as och
from tests.cases import make_case, oracle_for

pytestmark = pytest.mark.gpu

RTOL = 1e-10
WAVES = {'kinematic': 0, 'diffusive': 1, 'dynamic': 2}


def field_err(got, want, floor=1e-9):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    assert np.array_equal(np.isnan(got), np.isnan(want)), 'NaN pattern differs'
    scale = np.nanmax(np.abs(want)) if want.size else 0.0
    if not np.isfinite(scale) or scale == 0.0:
        return float(np.nanmax(np.abs(got - want))) if want.size else 0.0
    den = np.maximum(np.abs(want), floor * scale)
    return float(np.nanmax(np.abs(got - want) / den))


def gpu_for(c, wave, dt, allow_packed=True):
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200.engine import Strip
    from topoflow_old_b200.routing import setup_channels
    s = Strip(c['nx'], c['ny'])
    s.d8_build(c['dem'], c['dx'], c['dy'], c['dd'])
    assert np.array_equal(s.d8_get(L.D8_CODE), c['d8']['d8_grid'])
    cp = lambda v: v.copy() if np.ndim(v) else float(v)
    setup_channels(s, wave, c['law'], dt, c['da'], c['slope'].copy(), cp(c['width']),
                   cp(c['angle']), cp(c['rough']), sinu=c['sinu'], d0=c['d0'],
                   outlets=c['outlets'], allow_packed=allow_packed)
    return s


def compare(s, ch, tag, tol=RTOL, u_floor=1e-9):
    from topoflow_old_b200 import _lib as L
    worst = {}
    for name, fid in (('vol', L.F_VOL), ('d', L.F_D), ('u', L.F_U), ('Q', L.F_Q)):
        worst[name] = field_err(s.get(fid), getattr(ch, name), u_floor if name == 'u' else 1e-9)
    assert max(worst.values()) <= tol, (tag, worst)
    return worst


def run_pair(c, wave, dt, n_steps, allow_packed=True, check_every=None, batch=1):
    from topoflow_old_b200 import _lib as L
    ch = oracle_for(c, wave, dt)
    s = gpu_for(c, wave, dt, allow_packed)
    try:
        R = (ch.P_rain + ch.SM + ch.GW + ch.MR) - (ch.ET + ch.IN)
        inc = s.set_R(R)
        vol_R = 0.0
        hyd_g, hyd_o = [], []
        done = 0
        while done < n_steps:
            nb = min(batch, n_steps - done)
            taps, _, bad = s.step(nb)
            assert bad[0] == -1 and bad[1] == -1
            for k in range(nb):
                ch.update()
                vol_R += inc
                hyd_o.append([ch.Q_outlet.copy(), ch.u_outlet.copy(), ch.d_outlet.copy(),
                              ch.f_outlet.copy()])
                hyd_g.append(taps[k, 0])
            done += nb
            if check_every and done % check_every == 0:
                compare(s, ch, 'step %d' % done)
        worst = compare(s, ch, 'final')
        hyd_g, hyd_o = np.array(hyd_g), np.array(hyd_o, dtype=np.float64)
        for j, nm in enumerate('Qudf'):
            assert field_err(hyd_g[:, j], hyd_o[:, j]) <= RTOL, ('hydrograph', nm)
        assert abs(vol_R - float(ch.vol_R)) <= 1e-12 * max(1.0, abs(float(ch.vol_R)))
        return s, ch, worst
    except Exception:
        s.close()
        raise


@pytest.mark.parametrize('wave', ['kinematic', 'diffusive', 'dynamic'])
@pytest.mark.parametrize('packed', [True, False])
def test_three_waves_match_oracle(wave, packed):
    c = make_case(96, 80, seed=11, angle_mode='classes')
    dt = 0.5 if wave == 'dynamic' else 2.0
    s, ch, worst = run_pair(c, WAVES[wave], dt, 300, allow_packed=packed, check_every=100)
    try:
        assert ch.Q.max() > 0.0
        from topoflow_old_b200 import _lib as L
        # derived fields, recomputed on demand
        for name, fid in (('f', L.F_F), ('Rh', L.F_RH), ('A_wet', L.F_A_WET),
                          ('P_wet', L.F_P_WET), ('tau', L.F_TAU), ('u_star', L.F_U_STAR),
                          ('froude', L.F_FROUDE), ('S_free', L.F_S_FREE), ('width', L.F_WIDTH)):
            if name == 'S_free' and wave == 'kinematic':
                # never updated by the kinematic wave: the component keeps the
                # host copy made at channels_base.py:889
                with pytest.raises(L.TopoflowB200Error, match='kinematic'):
                    s.get(fid)
                continue
            assert field_err(s.get(fid), getattr(ch, name)) <= RTOL, name
        mm = s.minmax()
        o = ch.mins_and_maxes()
        want = [o['Q_min'], o['Q_max'], o['u_min'], o['u_max'], o['d_min'], o['d_max']]
        assert field_err(mm, np.array(want)) <= RTOL
    finally:
        s.close()


def test_first_step_is_bit_exact():
    """Before any cube root has fed back, vol and d carry only correctly
    rounded operations: they must match bit for bit."""
    from topoflow_old_b200 import _lib as L
    c = make_case(70, 50, seed=3, angle_mode='uniform', d0=0.01)
    ch = oracle_for(c, 0, 2.0)
    s = gpu_for(c, 0, 2.0)
    try:
        s.set_R(float(c['P_rain']))
        s.step(1)
        ch.update()
        assert np.array_equal(s.get(L.F_VOL), ch.vol)
        assert np.array_equal(s.get(L.F_D), ch.d)
    finally:
        s.close()


@pytest.mark.parametrize('law', [0, 1])
@pytest.mark.parametrize('mode', ['zero', 'uniform'])
def test_angles_and_friction_laws(law, mode):
    c = make_case(64, 64, seed=5 + law, angle_mode=mode, law=law)
    s, ch, _ = run_pair(c, 0, 2.0, 200)
    s.close()


def test_scalar_parameters_like_the_plane_case():
    c = make_case(48, 40, seed=9, scalar_params=True)
    for wave, dt in ((0, 2.0), (1, 2.0), (2, 0.5)):
        s, ch, _ = run_pair(c, wave, dt, 150)
        s.close()


def test_generic_layout_when_values_are_not_float32():
    from topoflow_old_b200 import _lib as L
    c = make_case(64, 48, seed=13, exact_f32=False, sinu=1.3, rain_grid=True)
    s, ch, _ = run_pair(c, 1, 2.0, 150)
    try:
        assert s.timing()[3] > 100       # float64 parameter rasters in use
    finally:
        s.close()


def test_overbank_rule_fires_like_the_reference():
    """5 m cells with channels up to 6 m wide: 'top_width > dw' fires for many
    cells on the very first step and for more as the depth grows; each such
    cell becomes a rectangle of width dw for good (channels_base.py:1449-1454).
    (The case is kept hydraulically stable, CFL < 1: with u*dt > ds the
    reference's own scheme amplifies 1-ulp differences without bound, so no
    two implementations of it agree there.)"""
    from topoflow_old_b200 import _lib as L
    c = make_case(64, 64, seed=21, cell=5.0, angle_mode='classes', rain_mmph=400.0)
    s, ch, _ = run_pair(c, 0, 0.5, 600, check_every=100, batch=25)
    try:
        fired = (ch.angle == 0) & (np.float64(c['angle']) != 0)
        assert fired.sum() > 300
        assert np.array_equal(s.get(L.F_WIDTH), ch.width)
        rule = (ch.width != c['width']) | fired
        assert np.array_equal(s.get(L.F_MUTATED)[rule], np.ones(int(rule.sum())))
    finally:
        s.close()


def test_batched_steps_equal_single_steps():
    c = make_case(80, 60, seed=17)
    s1, ch1, _ = run_pair(c, 0, 2.0, 120, batch=1)
    s2, ch2, _ = run_pair(c, 0, 2.0, 120, batch=40)
    from topoflow_old_b200 import _lib as L
    try:
        for f in (L.F_VOL, L.F_D, L.F_U, L.F_Q):
            assert np.array_equal(s1.get(f), s2.get(f))
    finally:
        s1.close()
        s2.close()


def test_ragged_and_tiny_grids():
    for nx, ny in ((3, 3), (5, 4), (33, 7), (129, 9)):
        c = make_case(nx, ny, seed=nx)
        s, ch, _ = run_pair(c, 0, 2.0, 30)
        s.close()


def test_sources_sinks_and_canals():
    from topoflow_old_b200 import _lib as L
    c = make_case(60, 50, seed=23)
    ch = oracle_for(c, 0, 2.0)
    s = gpu_for(c, 0, 2.0)
    try:
        src = ([10, 20], [12, 30], [0.5, 0.25])
        snk = ([30], [31], [0.01])
        cin, cout = ([25, 26], [20, 21]), ([40, 41], [22, 23])
        frac, outQ = [0.5, 1.0], [0.02, 0.0]
        ch.set_sources(*src)
        ch.set_sinks(*snk)
        ch.set_canals(cin[0], cin[1], cout[0], cout[1], frac, outQ)
        s.set_points(L.PT_SOURCES, *src)
        s.set_points(L.PT_SINKS, *snk)
        s.set_points(L.PT_CANALS_IN, cin[0], cin[1], frac)
        s.set_points(L.PT_CANALS_OUT, cout[0], cout[1], outQ)
        s.set_R(float(c['P_rain']))
        for n in range(60):
            _, cg, bad = s.step(1)
            ch.update()
            assert field_err(cg[0], ch.canals_in_Q) <= RTOL
        compare(s, ch, 'diversions')
    finally:
        s.close()


def test_bad_values_are_flagged_like_check_flow():
    """check_flow_depth / check_flow_velocity (channels_base.py:1982-2080):
    the first offending cell in row-major order is reported."""
    from topoflow_old_b200 import _lib as L
    c = make_case(40, 30, seed=29)
    # (a) NaN forcing -> NaN volume -> NaN depth everywhere water can be
    ch = oracle_for(c, 0, 2.0)
    s = gpu_for(c, 0, 2.0)
    try:
        ch.P_rain = och.scalar(np.nan)
        s.set_R(float('nan'))
        _, _, bad = s.step(1)
        with pytest.raises(RuntimeError, match='Negative depth found.'):
            ch.update()
        first = np.flatnonzero(~np.isfinite(ch.d.ravel()) | (ch.d.ravel() < 0))[0]
        assert bad[0] == first
    finally:
        s.close()
    # (b) a negative bed slope under the diffusive wave (no remove_bad_slopes):
    # S_free is clipped at 0 there, u = 0, nothing is flagged -- in both
    ch = oracle_for(c, 1, 2.0)
    ch.S_bed[12, 17] = -0.5
    c2 = dict(c)
    c2['slope'] = c['slope'].copy()
    c2['slope'][12, 17] = -0.5
    s = gpu_for(c2, 1, 2.0)
    try:
        s.set_R(float(c['P_rain']))
        for _ in range(20):
            _, _, bad = s.step(1)
            ch.update()
        assert bad[0] == -1 and bad[1] == -1
        compare(s, ch, 'negative slope, diffusive')
    finally:
        s.close()


def test_output_frames_are_the_float32_grids_the_reference_writes():
    """tf_chan_frame: np.float32(grid) (+ byteswap) done on the device, the copy
    overlapping later steps (rts_files.py:313-396)."""
    from topoflow_old_b200 import _lib as L
    c = make_case(157, 83, seed=31)                  # ragged: nx not a multiple of 32
    s = gpu_for(c, 1, 2.0)
    try:
        s.set_R(float(c['P_rain']))
        s.step(40)
        for fid in (L.F_Q, L.F_D, L.F_U, L.F_F, L.F_RH):
            want = np.float32(s.get(fid))
            got = s.frame(fid)
            assert got.dtype == np.float32 and got.shape == (83, 157)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), fid
            swapped = s.frame(fid, swap_endian=True)
            assert np.array_equal(swapped.view(np.uint32), want.byteswap().view(np.uint32)), fid
        assert np.float32(s.get(L.F_Q)).max() > 0
        # asynchronous: the frame is the state at the call, whatever runs afterwards
        buf_q, buf_d = s.pinned_frame(), s.pinned_frame()
        want_q, want_d = np.float32(s.get(L.F_Q)), np.float32(s.get(L.F_D))
        s.frame(L.F_Q, out=buf_q, wait=False)
        s.frame(L.F_D, swap_endian=True, out=buf_d, wait=False)   # waits on the device for the first copy
        s.step(7)
        s.frame_wait()
        assert np.array_equal(buf_q.view(np.uint32), want_q.view(np.uint32))
        assert np.array_equal(buf_d.view(np.uint32), want_d.byteswap().view(np.uint32))
        assert not np.array_equal(np.float32(s.get(L.F_Q)), want_q)   # the state did move on
    finally:
        s.close()


def test_update_R_on_the_device():
    """tf_chan_set_forcing / tf_chan_assemble_R against update_R and
    update_R_integral (channels_base.py:1103-1160) with a mix of grid and
    scalar fluxes."""
    from topoflow_old_b200 import _lib as L
    c = make_case(97, 61, seed=17)
    ch = oracle_for(c, 0, 2.0)
    s = gpu_for(c, 0, 2.0)
    rng = np.random.default_rng(5)
    try:
        grid = lambda scale: scale * rng.random((61, 97))
        ch.P_rain, ch.SM, ch.GW = grid(3e-5), och.scalar(2e-6), grid(1e-6)
        ch.MR, ch.ET, ch.IN = och.scalar(0.0), grid(5e-7), grid(4e-6)
        vol_R = 0.0
        for n in range(30):
            if n == 10:                                   # a flux changes type mid-run
                ch.SM = grid(1e-6)
            if n == 20:
                ch.IN = och.scalar(1e-6)
            for k, v in enumerate((ch.P_rain, ch.SM, ch.GW, ch.MR, ch.ET, ch.IN)):
                s.set_forcing(k, v)
            vol_R += s.assemble_R()
            want = (ch.P_rain + ch.SM + ch.GW + ch.MR) - (ch.ET + ch.IN)
            if n in (0, 10, 20):
                got = s.get(L.F_R)
                assert np.array_equal(got.view(np.uint64), want.view(np.uint64))     # bit for bit
            _, _, bad = s.step(1)
            ch.update()
        assert abs(vol_R - float(ch.vol_R)) <= 1e-12 * abs(float(ch.vol_R))
        compare(s, ch, 'device R')
        # all six scalar: R goes back to being a kernel argument
        for k in range(6):
            s.set_forcing(k, 1e-6 * (k + 1))
        inc = s.assemble_R()
        r = (1e-6 + 2e-6 + 3e-6 + 4e-6) - (5e-6 + 6e-6)
        assert inc == ((r * float(c['da'])) * 2.0) * (97 * 61)
        with pytest.raises(RuntimeError, match='R is a scalar'):
            s.get(L.F_R)
    finally:
        s.close()


@pytest.mark.parametrize('wave', ['kinematic', 'dynamic'])
def test_raster_lengths_and_no_lookahead_paths(wave, monkeypatch):
    """The fallbacks behind two layout decisions: ds / dw read from their
    rasters (taken when they are not a function of the code and the row, e.g.
    a sinuosity grid) and the L2 look-ahead switched off."""
    monkeypatch.setenv('TF_B200_NO_LEN_ROWS', '1')
    monkeypatch.setenv('TF_B200_PF_ROWS', '0')
    c = make_case(131, 77, seed=23)
    s, ch, _ = run_pair(c, WAVES[wave], 2.0 if wave == 'kinematic' else 0.5, 60)
    s.close()


def test_sinuosity_grid_takes_the_raster_path():
    c = make_case(90, 70, seed=29)
    rng = np.random.default_rng(3)
    c['sinu'] = 1.0 + 0.5 * np.float64(np.float32(rng.random((70, 90))))
    s, ch, _ = run_pair(c, 0, 2.0, 60)
    assert s.timing()[3] >= 72.0            # ds and dw come from their rasters
    s.close()
