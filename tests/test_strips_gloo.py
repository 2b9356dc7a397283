"""The N > 1 path on CPU: two processes (gloo), each owning a row strip plus
one halo row per side, run the routing oracle on their strips and trade halo
rows each step exactly as topoflow_old_b200.strips.halo_plan prescribes.  The
owned rows must equal the single-process run bit for bit -- i.e. one halo row
of the planned rasters per step is sufficient and placed at the right point
of the step."""
import os
import socket

import numpy as np
import pytest

from oracle import channels as och
from oracle import d8 as od8
from topoflow_old_b200 import strips
from tests.cases import make_case, oracle_for

N_STEPS = 40


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_strip_rows_partition():
    for ny, n in ((100, 3), (16384, 8), (7, 2), (9, 3)):
        rows = [strips.strip_rows(ny, n, r) for r in range(n)]
        assert rows[0][0] == 0 and sum(c for _, c in rows) == ny
        for (a, ca), (b, _) in zip(rows, rows[1:]):
            assert a + ca == b
        assert max(c for _, c in rows) - min(c for _, c in rows) <= 1
    with pytest.raises(ValueError):
        strips.strip_rows(5, 2, 0)
    assert strips.neighbours(4, 0) == (None, 1) and strips.neighbours(4, 3) == (2, None)
    assert strips.halo_plan('dynamic')[-1] == ('end', ['Qnext', 'u'])


def _local_oracle(case, wave, dt, r0, n):
    """Oracle on rows [r0-1, r0+n+1) clipped to the grid: owned rows + halos."""
    ny = case['ny']
    lo, hi = max(r0 - 1, 0), min(r0 + n + 1, ny)
    full = case['d8']['d8_grid']
    code = full[lo:hi].copy()
    pid = od8.parent_id_grid(np.pad(code, ((0, 0), (0, 0))))
    # parents are derived locally from the codes; a halo row's cells may point
    # outside the local window: drop those links (their results are discarded)
    pid_local = np.zeros_like(pid)
    nx = case['nx']
    rr, cc = np.divmod(np.arange(code.size), nx)
    for k, (dr, dc) in enumerate(od8.NBR):
        m = (code.ravel() == (1 << k))
        pr, pc = rr[m] + dr, cc[m] + dc
        ok = (pr >= 0) & (pr < code.shape[0]) & (pc >= 0) & (pc < nx)
        idx = np.flatnonzero(m)[ok]
        pid_local.ravel()[idx] = pr[ok] * nx + pc[ok]
        code.ravel()[np.flatnonzero(m)[~ok]] = 0
    b = dict(d8_grid=code, parent_ID_grid=pid_local, parent_IDs=od8.parent_ids(pid_local),
             ds=case['d8']['ds'][lo:hi], dw=case['d8']['dw'][lo:hi], noflow_IDs=od8.noflow_ids(code))
    b['w'], b['p'] = od8.flow_from_to(code, pid_local)
    sl = lambda v: v[lo:hi].copy() if np.ndim(v) else och.scalar(float(v))
    ch = och.ChannelsOracle(b, wave, dt=dt, da=case['da'], slope=sl(case['slope']), width=sl(case['width']),
                            angle=sl(case['angle']), nval=sl(case['rough']), outlet_ID=(1, 1))
    ch.P_rain = case['P_rain']
    return ch, lo, hi


def _worker(rank, world, port, wave_name, out_dir):
    import torch.distributed as dist
    import torch
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    wave = {'kinematic': 0, 'diffusive': 1, 'dynamic': 2}[wave_name]
    dt = 0.5 if wave == 2 else 2.0
    case = make_case(40, 36, seed=31)
    r0, n = strips.strip_rows(case['ny'], world, rank)
    ch, lo, hi = _local_oracle(case, wave, dt, r0, n)
    north, south = strips.neighbours(world, rank)
    first, last = r0 - lo, r0 - lo + n - 1        # local indices of the owned edge rows

    def exchange(arrays):
        for a in arrays:
            reqs, bufs = [], []
            if north is not None:
                reqs.append(dist.isend(torch.from_numpy(a[first].copy()), north))
                buf = torch.empty(a.shape[1], dtype=torch.float64)
                reqs.append(dist.irecv(buf, north))
                bufs.append((first - 1, buf))
            if south is not None:
                reqs.append(dist.isend(torch.from_numpy(a[last].copy()), south))
                buf = torch.empty(a.shape[1], dtype=torch.float64)
                reqs.append(dist.irecv(buf, south))
                bufs.append((last + 1, buf))
            for r in reqs:
                r.wait()
            for row, buf in bufs:
                a[row] = buf.numpy()

    plan = dict(strips.halo_plan(wave_name))
    for _ in range(N_STEPS):
        # one reference step, cut where the plan says rasters must cross
        ch.update_R(); ch.update_R_integral(); ch.update_discharge(); ch.update_diversions()
        ch.update_flow_volume(); ch.update_flow_depth()
        if 'after_depth' in plan:
            exchange([ch.d])
            ch.d_is_pos = (ch.d > 0); ch.d_is_neg = np.invert(ch.d_is_pos)
        ch.update_trapezoid_Rh()
        if not ch.KINEMATIC_WAVE:
            ch.update_free_surface_slope()
        ch.update_shear_stress(); ch.update_shear_speed(); ch.update_friction_factor()
        ch.update_velocity()
        ch.u[ch.d8['noflow_IDs']] = 0.0
        # "Qnext" = u * A_wet: the halo rows of its two factors
        exchange([ch.u, ch.A_wet])
        ch.time += ch.dt
    np.save(os.path.join(out_dir, 'strip_%d.npy' % rank),
            np.stack([ch.vol[first:last + 1], ch.d[first:last + 1], ch.u[first:last + 1]]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize('wave_name', ['kinematic', 'diffusive', 'dynamic'])
def test_two_strips_equal_one_grid(tmp_path, wave_name):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), wave_name, str(tmp_path)), nprocs=world, join=True)
    wave = {'kinematic': 0, 'diffusive': 1, 'dynamic': 2}[wave_name]
    case = make_case(40, 36, seed=31)
    ch = oracle_for(case, wave, 0.5 if wave == 2 else 2.0)
    for _ in range(N_STEPS):
        ch.update()
    got = np.concatenate([np.load(str(tmp_path / ('strip_%d.npy' % r))) for r in range(world)], axis=1)
    assert ch.Q.max() > 0
    for k, name in enumerate(('vol', 'd', 'u')):
        assert np.array_equal(got[k], getattr(ch, name)), name


# ---------------------------------------------------------------------------
# contributing area on strips: the owner-computes rounds of csrc/tf_d8.cu
# (area_strips), restated in NumPy and run by two gloo processes
# ---------------------------------------------------------------------------
def _area_worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    case = make_case(40, 36, seed=31)
    code = case['d8']['d8_grid']
    ny, nx = code.shape
    r0, n = strips.strip_rows(ny, world, rank)
    north, south = strips.neighbours(world, rank)
    pa = np.float32(np.float64(case['da']) / 1e6)
    # local rows 1..n are owned, row 0 and row n+1 are the neighbours' edge rows
    C = np.zeros((n + 2, nx), dtype=np.uint8)
    C[1:n + 1] = code[r0:r0 + n]
    if north is not None:
        C[0] = code[r0 - 1]
    if south is not None:
        C[n + 1] = code[r0 + n]
    A = np.zeros((n + 2, nx), dtype=np.float32)
    kcode = {c: k for k, c in enumerate(od8.CODE_LIST)}

    def children(y, x):
        """(k, yy, xx) of the neighbours that drain into (y, x), in the reference's order k = 0..7."""
        out = []
        for k, (dr, dc) in enumerate(od8.NBR):
            yy, xx = y + dr, x + dc
            if 0 <= yy < n + 2 and 0 <= xx < nx and C[yy, xx] == od8.CODE_LIST[(k + 4) % 8]:
                out.append((k, yy, xx))
        return out

    pending = {(y, x): len(children(y, x)) for y in range(1, n + 1) for x in range(nx) if C[y, x]}
    credited = set()

    def climb(y, x):
        while True:
            acc = np.float32(pa)
            for _, yy, xx in children(y, x):
                acc = np.float32(acc + A[yy, xx])
            A[y, x] = acc
            dr, dc = od8.NBR[kcode[int(C[y, x])]]
            py, px = y + dr, x + dc
            if not (1 <= py <= n) or not (0 <= px < nx) or C[py, px] == 0:
                return                                   # the neighbour strip (or nobody) goes on
            pending[(py, px)] -= 1
            if pending[(py, px)]:
                return
            y, x = py, px

    for (y, x), p in list(pending.items()):
        if p == 0:
            climb(y, x)
    rounds = 0
    while True:
        rounds += 1
        for peer, mine, theirs in ((north, 1, 0), (south, n, n + 1)):
            if peer is None:
                continue
            req = dist.isend(torch.from_numpy(A[mine].copy()), peer)
            buf = torch.empty(nx, dtype=torch.float32)
            dist.recv(buf, peer)
            req.wait()
            A[theirs] = buf.numpy()
        n_new = 0
        for y, edge in ((1, 0), (n, n + 1)):
            for x in range(nx):
                if not C[y, x]:
                    continue
                for k, yy, xx in children(y, x):
                    if yy == edge and A[yy, xx] > 0 and (y, x, k) not in credited:
                        credited.add((y, x, k))
                        n_new += 1
                        pending[(y, x)] -= 1
                        if pending[(y, x)] == 0:
                            climb(y, x)
        t = torch.tensor([n_new])
        dist.all_reduce(t)
        if int(t.item()) == 0:
            break
    np.save(os.path.join(out_dir, 'area_%d.npy' % rank), A[1:n + 1])
    np.save(os.path.join(out_dir, 'rounds_%d.npy' % rank), np.array([rounds]))
    dist.barrier()
    dist.destroy_process_group()


def test_area_rounds_on_two_strips_equal_one_grid(tmp_path):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_area_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    case = make_case(40, 36, seed=31)
    want = od8.area_grid_fast(case['d8']['d8_grid'], case['da'], 'km^2')
    got = np.concatenate([np.load(str(tmp_path / ('area_%d.npy' % r))) for r in range(world)], axis=0)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert int(np.load(str(tmp_path / 'rounds_0.npy'))[0]) >= 2      # at least one flow path crossed the cut
