#!/usr/bin/env python
"""Headline benchmark: grid cell-updates/s of the fp64 channel-routing step.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      (the CPU path, NumPy, 1 core)

Workload (BASELINE.json north_star): kinematic-wave, Manning friction, fp64
state on a synthetic fractal DEM of --size x --size cells per GPU (default
16384 x 16384), uniform rain pulse, D8 grid built on the GPU first.  With N
GPUs the grid is N strips of --size rows stacked north to south (weak
scaling), one halo row of Q exchanged per step over NCCL.

One JSON line on stdout (rank 0); everything else goes to stderr.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

_OUT = sys.stdout
ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ALGO_BYTES = {'kinematic': 97.0, 'diffusive': 113.0, 'dynamic': 129.0}   # SURVEY.md 8(d), R a grid
ALGO_BYTES_SCALAR_R = 8.0          # ... minus the R read when R is a 0-d scalar (89 / 105 / 121)
WAVE_ID = {'kinematic': 0, 'diffusive': 1, 'dynamic': 2}
DT = {'kinematic': 2.0, 'diffusive': 2.0, 'dynamic': 0.5}
CELL = 30.0
RAIN_MPS = 50.0 / 3.6e6            # 50 mm/h uniform rain pulse


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        return float(json.load(open(p))['hbm_gbs']), 'measured (MEASURED_PEAKS.json)'
    return 6650.0, 'fallback (B200_PROFILING.md)'


# ---------------------------------------------------------------------------
# inputs
# ---------------------------------------------------------------------------
def synth_inputs(nx, global_ny, row0, nrows, seed=20170226, n_procs=1):
    """DEM strip (+ its halo rows) and the float32 parameter grids, as the
    reference would read them from RTG files."""
    from concurrent.futures import ThreadPoolExecutor
    from topoflow_old_b200.utils import synth
    lo, hi = max(row0 - 1, 0), min(row0 + nrows + 1, global_ny)
    chunks = [(a, min(a + 512, hi)) for a in range(lo, hi, 512)]
    with ThreadPoolExecutor(max_workers=max(2, min(32, (os.cpu_count() or 1) // n_procs))) as ex:
        parts = list(ex.map(lambda ab: synth.fractal_dem(nx, global_ny, row0=ab[0], nrows=ab[1] - ab[0],
                                                         seed=seed, cell=CELL, roughness=8.0), chunks))
    dem = np.concatenate(parts, axis=0)
    north = dem[0] if row0 > 0 else None
    south = dem[-1] if row0 + nrows < global_ny else None
    dem = dem[(row0 - lo):(row0 - lo) + nrows]
    nval = synth.uniform_field(nx, global_ny, 0.03, 0.15, row0=row0, nrows=nrows, seed=seed + 3)
    width = synth.uniform_field(nx, global_ny, 1.0, 6.0, row0=row0, nrows=nrows, seed=seed + 1)
    angle = synth.class_field(nx, global_ny, [20.0, 30.0, 45.0], row0=row0, nrows=nrows,
                              seed=seed + 2, block=64)
    return dem, north, south, nval, width, angle


def workload_text(args, nranks):
    return 'synthetic %dx%d fractal DEM %s, %s-wave channels, Manning, uniform rain pulse (0-d R), D8 built on GPU' % (
        args.size, args.size, 'per GPU' if args.scaling == 'weak' else 'cut into %d row strips' % nranks, args.wave)


def global_rows(args, nranks):
    """Rows of the global grid: weak scaling gives every GPU size x size cells,
    strong scaling cuts one size x size grid into nranks strips."""
    return args.size * nranks if args.scaling == 'weak' else args.size


def build_strip(args, rank, nranks, device, unique_id=None, reduce_min=None, nx=None, global_ny=None,
                wave=None, want_area=False, quiet=False, keep_dem=False):
    """D8 grid and routing state of this rank's strip, ready to step."""
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200.engine import Strip
    from topoflow_old_b200.routing import setup_channels
    from topoflow_old_b200.utils import synth
    from topoflow_old_b200 import strips
    wave = wave or args.wave
    nx = nx or args.size
    global_ny = global_ny or global_rows(args, nranks)
    row0, ny = strips.strip_rows(global_ny, nranks, rank)
    log_ = (lambda *a: None) if quiet else log
    t0 = time.time()
    dem, north, south, nval, width, angle = synth_inputs(nx, global_ny, row0, ny, n_procs=nranks)
    log_('[rank %d] synthetic inputs %.1fs' % (rank, time.time() - t0))
    s = Strip(nx, ny, device=device)
    if nranks > 1:
        s.set_strip(row0, global_ny, rank > 0, rank < nranks - 1)
        s.comm_init(rank, nranks, unique_id)
    dx = np.full(ny, CELL, dtype=np.float32)
    dd = np.full(ny, np.float32(np.sqrt(np.float64(CELL) ** 2 * 2)), dtype=np.float32)
    s.pits_filled = None
    if getattr(args, 'fill_pits', False) and nranks == 1:
        # the generator may fill the DEM's depressions first (not timed, not the default:
        # the shipped cfgs run with FILL_PITS_IN_Z0 = 0); the fill is the library's own tf_fill_pits
        t0 = time.time()
        dem, n_raised, sweeps = s.fill_pits(dem)
        s.pits_filled = {'cells_raised': n_raised, 'sweeps': sweeps, 'seconds': round(time.time() - t0, 2)}
        log_('[rank %d] pit filling %.2fs %s' % (rank, time.time() - t0, s.pits_filled))
    t0 = time.time()
    stats = s.d8_build(dem, dx, dx, dd)
    log_('[rank %d] D8 build %.2fs %s' % (rank, time.time() - t0, stats))
    s.d8_stats = stats
    code = s.d8_get(L.D8_CODE)
    ds = s.d8_get(L.D8_DS)
    dw = s.d8_get(L.D8_DW)
    slope = synth.d8_slope(dem, code, ds, north=north, south=south)
    s.dem_host = dem if keep_dem else None
    del dem
    width = np.float64(width)
    width[width == 0] = dw[width == 0]                       # channels_base.py:2182-2183
    angle = np.float64(angle)
    angle *= np.pi / 180.0                                   # :2191
    outlets = [(ny // 2, nx // 2), (ny - 2, nx - 2)]
    t0 = time.time()
    if want_area:
        t1 = time.time()
        s.area_unresolved = s.d8_area(CELL * CELL, units_km=True)
        s.area_seconds = time.time() - t1
    s.noflow_fraction = float((code == 0).mean())
    setup_channels(s, WAVE_ID[wave], L.MANNING, DT[wave], CELL * CELL, np.float64(slope),
                   width, angle, np.float64(nval), sinu=1.0, d0=0.0, outlets=outlets,
                   allow_packed=not args.generic, ds=ds, reduce_min=reduce_min)
    log_('[rank %d] routing setup %.1fs' % (rank, time.time() - t0))
    s.set_R(RAIN_MPS)
    s.host_inputs = dict(slope=np.float64(slope), width=width, angle=angle, nval=np.float64(nval),
                         code=code, ds=ds) if quiet else None
    return s


# ---------------------------------------------------------------------------
# parity before timing: strips == one strip (bit for bit) == oracle (1e-10)
# ---------------------------------------------------------------------------
PARITY_NX, PARITY_ROWS_PER_RANK, PARITY_STEPS = 2048, 256, 8


def _field_err(got, want, floor=1e-9):
    """max |got - want| / max(|want|, floor * max|want|), as tests/test_gpu_channels.py."""
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    if not np.array_equal(np.isnan(got), np.isnan(want)):
        return float('inf')
    scale = np.nanmax(np.abs(want)) if want.size else 0.0
    if not np.isfinite(scale) or scale == 0.0:
        return float(np.nanmax(np.abs(got - want))) if want.size else 0.0
    return float(np.nanmax(np.abs(got - want) / np.maximum(np.abs(want), floor * scale)))


def parity_leg(args, D):
    """A 2048 x (256 N) grid routed three ways, for all three waves: cut into the
    N strips of this run (halo rows over NCCL), as one strip on rank 0, and by
    the NumPy oracle on rank 0.  Strips and the single strip must agree bit for
    bit (D8 codes, contributing area, vol, d, u, Q); the device and the oracle
    within 1e-10 (codes and area exactly).  Returns the dict for the JSON line
    (rank 0) -- the caller exits non-zero when `ok` is false."""
    from topoflow_old_b200 import _lib as L
    from topoflow_old_b200.engine import Strip
    nx, gny = PARITY_NX, PARITY_ROWS_PER_RANK * D.world
    fields = (('vol', L.F_VOL), ('d', L.F_D), ('u', L.F_U), ('Q', L.F_Q))
    out = {'grid': [gny, nx], 'steps': PARITY_STEPS, 'strips': D.world, 'waves': {}}
    strips_equal, worst, codes_ok, area_ok = True, 0.0, True, True
    oracle_d8 = None
    t_start = time.time()
    for wave in ('kinematic', 'diffusive', 'dynamic'):
        uid = None
        if D.world > 1:
            uid = Strip.comm_unique_id() if D.rank == 0 else np.zeros(128, dtype=np.uint8)
            uid = D.bcast_bytes(uid)
        s = build_strip(args, D.rank, D.world, D.local_rank, uid,
                        reduce_min=(lambda v: -D.max(-v)) if D.world > 1 else None,
                        nx=nx, global_ny=gny, wave=wave, want_area=True, quiet=True)
        _, _, bad = s.step(PARITY_STEPS)
        mine = {n: s.get(f) for n, f in fields}
        mine['code'] = s.d8_get(L.D8_CODE)
        mine['area'] = s.d8_get(L.D8_AREA)
        mine['bad'] = bad.copy()
        s.close()
        parts = D.gather(mine)
        if D.rank != 0:
            continue
        whole = {k: np.concatenate([p[k] for p in parts], axis=0) for k in mine if k != 'bad'}
        no_bad = all(int(p['bad'][0]) == -1 and int(p['bad'][1]) == -1 for p in parts)
        one = build_strip(args, 0, 1, D.local_rank, None, None, nx=nx, global_ny=gny, wave=wave,
                          want_area=True, quiet=True)
        one.step(PARITY_STEPS)
        single = {n: one.get(f) for n, f in fields}
        single['code'] = one.d8_get(L.D8_CODE)
        single['area'] = one.d8_get(L.D8_AREA)
        inp = one.host_inputs
        one.close()
        eq = all(np.array_equal(whole[k], single[k]) for k in single)
        strips_equal = strips_equal and eq and no_bad
        # ---- the oracle on the same inputs (rank 0, host) ----
        from oracle import channels as och
        from oracle import d8 as od8
        if oracle_d8 is None:
            dem = synth_inputs(nx, gny, 0, gny)[0]
            oracle_d8 = od8.build(dem, CELL, CELL)
            oracle_d8['A'] = od8.area_grid_fast(oracle_d8['d8_grid'], np.float64(CELL * CELL), 'km^2')
        codes_ok = codes_ok and np.array_equal(single['code'], oracle_d8['d8_grid'])
        area_ok = area_ok and np.array_equal(single['area'].view(np.uint32), oracle_d8['A'].view(np.uint32))
        ch = och.ChannelsOracle(oracle_d8, WAVE_ID[wave], dt=DT[wave], da=np.float64(CELL * CELL),
                                slope=inp['slope'].copy(), width=np.float64(inp['width']).copy(),
                                angle=np.float64(inp['angle']).copy(), nval=inp['nval'].copy(),
                                outlet_ID=(gny // 2, nx // 2))
        ch.P_rain = och.scalar(RAIN_MPS)
        for _ in range(PARITY_STEPS):
            ch.update()
        errs = {n: _field_err(single[n], getattr(ch, n)) for n, _ in fields}
        worst = max(worst, max(errs.values()))
        out['waves'][wave] = {'strips_bit_identical': bool(eq and no_bad), 'oracle_rel_err': max(errs.values())}
    if D.rank != 0:
        return None
    out.update(strips_bit_identical=bool(strips_equal), oracle_rel_err=worst,
               d8_codes_equal_oracle=bool(codes_ok), area_equal_oracle=bool(area_ok),
               tolerance=1e-10, seconds=round(time.time() - t_start, 1))
    out['ok'] = bool(strips_equal and codes_ok and area_ok and worst <= 1e-10)
    return out


# ---------------------------------------------------------------------------
# clocks during the timed region
# ---------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock and throttle reasons while the timed region runs (NVML in
    process, every 5 ms; the nvidia-smi line of B200_PROFILING.md is the
    fallback)."""
    REASONS = (('hw_slowdown', 0x8), ('sw_power_cap', 0x4), ('hw_thermal_slowdown', 0x40),
               ('sw_thermal_slowdown', 0x20))

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.sm, self.mask, self.stop_flag = device, [], 0, False
        self.sm_max, self.nvml, self.h = None, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get('CUDA_VISIBLE_DEVICES')
            idx = int(vis.split(',')[device]) if vis and vis.split(',')[device].isdigit() else device
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception as e:                      # noqa: BLE001
            log('NVML unavailable (%s); falling back to nvidia-smi' % e)

    def _sample(self):
        if self.nvml is not None:
            n = self.nvml
            self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
            self.mask |= int(n.nvmlDeviceGetCurrentClocksEventReasons(self.h))
            time.sleep(0.005)
            return
        q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_power_cap,'
             'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown')
        out = subprocess.run(['nvidia-smi', '-i', str(self.device), '--query-gpu=' + q,
                              '--format=csv,noheader,nounits'], capture_output=True, text=True,
                             timeout=5).stdout.strip()
        if out:
            r = [x.strip() for x in out.split(',')]
            self.sm.append(float(r[0]))
            self.sm_max = float(r[1])
            for i, (_, bit) in enumerate(self.REASONS):
                if r[2 + i] == 'Active':
                    self.mask |= bit

    def run(self):
        while not self.stop_flag:
            try:
                self._sample()
            except Exception:                       # noqa: BLE001
                time.sleep(0.05)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.sm:
            return {'sm_mhz': None, 'sm_max_mhz': self.sm_max, 'reasons': ['no sample taken']}
        sm = sorted(self.sm)
        return {'sm_mhz': sm[len(sm) // 2], 'sm_min_mhz': sm[0], 'sm_max_mhz': self.sm_max,
                'reasons': [n for n, bit in self.REASONS if self.mask & bit], 'samples': len(sm)}


# ---------------------------------------------------------------------------
# CPU path (the oracle: NumPy restatement of the reference, one core)
# ---------------------------------------------------------------------------
def cpu_path(wave, size, steps, warmup):
    """cell-updates/s of the reference algorithm on the host, on a size x size
    sample of the same synthetic workload."""
    from oracle import channels as och
    from oracle import d8 as od8
    from topoflow_old_b200.utils import synth
    dem = synth.fractal_dem(size, size, cell=CELL, roughness=8.0)
    b = od8.build(dem, CELL, CELL)
    slope = synth.d8_slope(dem, b['d8_grid'], b['ds'])
    nval = synth.uniform_field(size, size, 0.03, 0.15, seed=20170229)
    width = synth.uniform_field(size, size, 1.0, 6.0, seed=20170227)
    angle = synth.class_field(size, size, [20.0, 30.0, 45.0], seed=20170228, block=64)
    inp = och.prepare_inputs(b, slope, width, angle, nval=nval)
    ch = och.ChannelsOracle(b, WAVE_ID[wave], dt=DT[wave], da=np.float64(CELL * CELL),
                            slope=inp['slope'], width=inp['width'], angle=inp['angle'],
                            nval=inp['nval'], outlet_ID=(size // 2, size // 2))
    ch.P_rain = och.scalar(RAIN_MPS)
    for _ in range(warmup):
        ch.update()
    t0 = time.perf_counter()
    for _ in range(steps):
        ch.update()
    el = time.perf_counter() - t0
    return size * size * steps / el, el


# ---------------------------------------------------------------------------
# distributed plumbing (torch.distributed only carries the NCCL id, the
# barriers and the max-over-ranks of the timings)
# ---------------------------------------------------------------------------
class Dist(object):
    def __init__(self, n_gpus):
        self.rank = int(os.environ.get('RANK', '0'))
        self.world = int(os.environ.get('WORLD_SIZE', '1'))
        self.local_rank = int(os.environ.get('LOCAL_RANK', '0'))
        self.torch = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            self.torch, self.dist = torch, dist
            os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
            torch.cuda.set_device(self.local_rank)
            dist.init_process_group(backend='nccl', device_id=torch.device('cuda', self.local_rank))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def bcast_bytes(self, arr):
        if self.world == 1:
            return arr
        t = self.torch.from_numpy(arr.copy()).cuda()
        self.dist.broadcast(t, src=0)
        return t.cpu().numpy()

    def max(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device='cuda')
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device='cuda')
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def gather(self, obj):
        """obj of every rank, in rank order (a list on every rank)."""
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def time_steps(s, D, K, W):
    """W untimed + K timed steps of a ready strip set; (max-over-ranks ms, launches)."""
    s.step_async(W)
    s.sync()
    s.timing()
    D.barrier()
    if D.world > 1:
        s.step_async(1)
        s.sync()
        s.timing()
    s.step_async(K)
    ms_dev, _, launches, _ = s.timing()
    s.sync()
    D.barrier()
    return D.max(ms_dev), launches


def strong_leg(args, D):
    """BASELINE.json configs[3] next to the headline, whenever several GPUs run:
    ONE 16384 x 16384 grid, dynamic wave, cut into N row strips (strong scaling),
    and the same grid on rank 0 alone for the efficiency."""
    import copy
    a = copy.copy(args)
    a.wave, a.scaling, a.size = 'dynamic', 'strong', args.strong_size
    uid = Strip_uid(D)
    s = build_strip(a, D.rank, D.world, D.local_rank, uid, reduce_min=lambda v: -D.max(-v), quiet=True)
    ms, launches = time_steps(s, D, args.steps, max(args.warmup, 3))
    _, _, bad = s.step(1)
    s.close()
    one_ms = None
    if D.rank == 0:
        one = build_strip(a, 0, 1, D.local_rank, None, None, quiet=True)
        one.step_async(max(args.warmup, 3)); one.sync(); one.timing()
        one.step_async(args.steps)
        one_ms = one.timing()[0]
        one.close()
    D.barrier()
    if D.rank != 0:
        return None
    cells = float(a.size) * a.size
    return {'workload': workload_text(a, D.world), 'scaling': 'strong', 'n_gpus': D.world,
            'ms_per_step': ms / args.steps, 'value': cells * args.steps / (ms * 1e-3), 'unit': 'cell-updates/s',
            'single_gpu_ms_per_step': one_ms / args.steps,
            'efficiency': (one_ms / ms) / D.world, 'bad_flags': [int(bad[0]), int(bad[1])]}


def Strip_uid(D):
    from topoflow_old_b200.engine import Strip
    if D.world == 1:
        return None
    uid = Strip.comm_unique_id() if D.rank == 0 else np.zeros(128, dtype=np.uint8)
    return D.bcast_bytes(uid)


def coupled_forcing(nx, global_ny, row0, nrows, dem_lo, dem_hi, dem):
    """Inputs of the provider chain on this strip (BASELINE.json configs[4]): rain and soil
    conductivity vary cell by cell, the air temperature falls with elevation (+6 .. -6 deg C,
    so the upper part of the grid holds snow), the radiation terms are uniform."""
    from topoflow_old_b200.utils import synth
    u = lambda seed: synth.uniform_field(nx, global_ny, 0.0, 1.0, row0=row0, nrows=nrows, seed=seed, dtype=np.float64)
    T_air = 6.0 - 12.0 * (np.float64(dem) - dem_lo) / max(dem_hi - dem_lo, 1e-9)
    met = dict(dt=60.0, P=(200.0 / 3.6e6) * (0.5 + u(51)), T_air=T_air, T_surf=T_air - 1.0, Qn_SW=300.0,
               Qn_LW=-50.0, rain_duration=1800.0)
    snow = dict(dt=60.0, c0=2.7, T0=-0.2, h0_snow=0.3 * u(52), h0_swe=0.1 * u(52), rho_snow=300.0)
    evap = dict(dt=60.0, alpha=1.2, K_soil=0.45, soil_x=0.05, T_soil_x=0.0)
    infil = dict(dt=60.0, Ks=7.2e-6 * (0.5 + u(53)), Ki=1e-7, qs=0.485, qi=0.375808, G=0.724)
    return met, snow, evap, infil


def run_coupled(args):
    """met + degree-day snow + Priestley-Taylor evaporation + Green-Ampt infiltration + kinematic
    routing under EMELI's loop, every provider on the device (topoflow_old_b200/coupled.py).  A step
    is one channel step (dt 2 s); the providers update every 60 s = 30 steps, so K timed steps
    hold K / 30 rounds of provider kernels and re-assemblies of R."""
    from topoflow_old_b200.coupled import CoupledRun
    D = Dist(args.gpus)
    args.wave = 'kinematic'
    parity = None
    if not args.no_parity:
        parity = parity_leg(args, D)
        ok = D.max(0.0 if (parity is None or parity['ok']) else 1.0) == 0.0
        if not ok:
            if D.rank == 0:
                print(json.dumps({'error': 'parity check failed', 'parity': parity}), file=_OUT, flush=True)
            D.close()
            sys.exit(3)
    uid = Strip_uid(D)
    s = build_strip(args, D.rank, D.world, D.local_rank, uid,
                    reduce_min=(lambda v: -D.max(-v)) if D.world > 1 else None, keep_dem=True)
    dem = s.dem_host
    lo, hi = -D.max(-float(dem.min())), D.max(float(dem.max()))
    global_ny = global_rows(args, D.world)
    from topoflow_old_b200 import strips
    row0, ny = strips.strip_rows(global_ny, D.world, D.rank)
    met, snow, evap, infil = coupled_forcing(s.nx, global_ny, row0, ny, lo, hi, dem)
    del dem
    s.dem_host = None
    t0 = time.time()
    run = CoupledRun(s, DT['kinematic'], met, snow, evap, infil)
    del met, snow, evap, infil
    log('[rank %d] providers on the device %.1fs' % (D.rank, time.time() - t0))
    K, W = args.steps, max(args.warmup, 3)
    dt = DT['kinematic']
    sampler = ClockSampler(D.local_rank)
    run.run_until(dt * (1 + W))                       # (the initial pass already made one step)
    s.sync(); s.timing()
    D.barrier()
    sampler.start()
    t0 = time.perf_counter()
    run.run_until(dt * (1 + W + K))
    ms_dev, n_steps, launches, layout_bytes = s.timing()
    s.sync()
    D.barrier()
    e2e_s = D.max(time.perf_counter() - t0)
    ms = D.max(ms_dev)
    clocks = sampler.summary()
    cells = float(s.nx) * s.ny
    total_cells = D.sum(cells)
    assert run.bad[0] == -1 and run.bad[1] == -1, run.bad
    peak, peak_src = peaks()
    line = None
    if D.rank == 0:
        algo = ALGO_BYTES['kinematic'] + 8.0          # grid R (97 B) + its rate of change R1
        line = {
            'metric': 'grid cell-updates/sec (fp64 routing step)', 'value': total_cells * K / (ms * 1e-3),
            'unit': 'cell-updates/s', 'n_gpus': D.world, 'steps': K, 'warmup': W, 'ms_per_step': ms / K,
            'higher_is_better': True, 'scaling': args.scaling, 'vs_baseline': None, 'dtype': 'f64',
            'data': 'synthetic',
            'config': {'workload': 'coupled: met + snow degree-day + evap Priestley-Taylor + infil Green-Ampt (every '
                                   '60 s, on the device) + kinematic routing (2 s), synthetic %dx%d fractal DEM %s' % (
                                       args.size, args.size,
                                       'per GPU' if args.scaling == 'weak' else 'cut into %d row strips' % D.world),
                       'grid_per_gpu': [int(s.ny), int(s.nx)], 'wave': 'kinematic',
                       'provider_updates_in_timed_region': {k: int(v[2]) for k, v in run.clock.items()},
                       'l2_policy': 'inputs larger than L2'},
            'e2e': {'value': total_cells * K / e2e_s, 'unit': 'cell-updates/s', 'h2d_bytes_per_step': 48,
                    'd2h_bytes_per_step': 8 * 4 * s.n_outlets + 16,
                    'note': 'host loop of CoupledRun.run_until: provider kernels, interpolation, R re-assembly and '
                            'the channel steps, outlet taps copied back; no grid crosses PCIe'},
            'gpu_launches': int(launches), 'clocks': clocks,
            'roofline': {'bound': 'hbm', 'achieved': algo * cells / (ms / K * 1e-3) / 1e9, 'peak': peak, 'unit': 'GB/s',
                         'frac': algo * cells / (ms / K * 1e-3) / 1e9 / peak, 'traffic': None,
                         'peak_source': peak_src, 'algorithmic_bytes_per_cell': algo,
                         'note': 'step kernel bytes only (grid R + R1); the provider kernels of a 60 s round are '
                                 'inside ms_per_step but not in the byte count'},
            'volumes_m3': run.vol,
            'parity': parity,
        }
    s.close()
    D.close()
    if line is not None:
        print(json.dumps(line), file=_OUT, flush=True)


def run_gpu(args):
    from topoflow_old_b200.engine import Strip
    D = Dist(args.gpus)
    if D.world != args.gpus:
        log('warning: --gpus %d but WORLD_SIZE %d; using WORLD_SIZE' % (args.gpus, D.world))
    uid = None
    if D.world > 1:
        uid = Strip.comm_unique_id() if D.rank == 0 else np.zeros(128, dtype=np.uint8)
        uid = D.bcast_bytes(uid)
    # ---- parity first: a run whose strips or whose arithmetic are wrong is not timed ----
    parity = None
    if not args.no_parity:
        parity = parity_leg(args, D)
        ok = D.max(0.0 if (parity is None or parity['ok']) else 1.0) == 0.0
        if D.rank == 0:
            log('parity: %s' % json.dumps(parity))
        if not ok:
            if D.rank == 0:
                print(json.dumps({'error': 'parity check failed', 'parity': parity}), file=_OUT, flush=True)
            D.close()
            sys.exit(3)
    s = build_strip(args, D.rank, D.world, D.local_rank, uid,
                    reduce_min=(lambda v: -D.max(-v)) if D.world > 1 else None,
                    want_area=args.with_area)
    # torch creates its own communicator lazily at the first collective: do that (and let every
    # rank finish its set-up) well before the timed region, or the skew lands inside it
    D.barrier()
    D.max(0.0)
    cells = float(s.nx) * s.ny                        # per GPU
    area_s = D.max(getattr(s, 'area_seconds', 0.0)) if args.with_area else None
    noflow = D.sum(s.noflow_fraction) / D.world
    K, W = args.steps, max(args.warmup, 0)

    # ---- device-resident throughput ("value") -------------------------------
    sampler = ClockSampler(D.local_rank)              # NVML start-up costs ms and differs per rank:
    s.step_async(W)                                   # keep it out of the timed region
    s.sync()
    s.timing()
    D.barrier()
    D.barrier()
    if D.world > 1:
        # one more untimed step: its halo exchange lines the ranks up to within a step, so no
        # rank's event window contains another rank's host-side start-up skew
        s.step_async(1)
        s.sync()
        s.timing()
    sampler.start()
    t0 = time.perf_counter()
    s.step_async(K)
    ms_dev, n_steps, launches, layout_bytes = s.timing()      # CUDA events on the stream
    s.sync()
    D.barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    log('[rank %d] %d steps: %.3f ms on the device (%.4f ms/step)' % (D.rank, K, ms_dev, ms_dev / K))
    ms = D.max(ms_dev)
    total_cells = D.sum(cells)
    value = total_cells * K / (ms * 1e-3)

    # ---- end to end: one public call per step, host buffers -------------------
    D.barrier()
    t0 = time.perf_counter()
    Ke = K
    for _ in range(Ke):
        s.set_R(RAIN_MPS, want_sum=False)             # this step's forcing (6 scalars -> R)
        taps, _, bad = s.step(1)                      # outlet Q,u,d,f + bad-value flags back
    D.barrier()
    e2e_s = D.max(time.perf_counter() - t0)
    e2e = total_cells * Ke / e2e_s
    clocks = sampler.summary()
    assert bad[0] == -1 and bad[1] == -1, 'bad depth / velocity flagged: %s' % bad

    # ---- the same with a rain GRID handed over every step (update_R on the device) ----
    e2e_grid = None
    if not args.no_grid_rain:
        from topoflow_old_b200 import _lib as L
        P = s.pinned_array((s.ny, s.nx), np.float64)
        P[:] = RAIN_MPS
        for k in range(1, 6):
            s.set_forcing(k, 0.0)
        Kg = max(2, min(K, 5))
        D.barrier()
        t0 = time.perf_counter()
        for _ in range(Kg):
            s.set_forcing(L.R_P_RAIN, P)              # 8 B/cell host -> device
            s.assemble_R(want_sum=True)               # R and the vol_R increment, on the device
            taps, _, bad = s.step(1)
        D.barrier()
        e2e_grid = total_cells * Kg / D.max(time.perf_counter() - t0)
        s.set_R(RAIN_MPS, want_sum=False)
    s.timing()
    strong = None
    if D.world > 1 and args.scaling == 'weak' and not args.no_strong_leg:
        s.sync()
        strong = strong_leg(args, D)

    line = None
    peak, peak_src = peaks()
    if D.rank == 0:
        # the step kernels of one time step (one fused kernel for the kinematic wave, phase A +
        # phase B for the other two), timed together with CUDA events on the launching stream
        kernel_ms = ms_dev / K
        algo = ALGO_BYTES[args.wave] - ALGO_BYTES_SCALAR_R        # the bench's R is a 0-d scalar
        achieved = algo * cells / (kernel_ms * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, 'profiles', 'traffic_%s.json' % args.wave)
        if os.path.exists(tp):
            tj = json.load(open(tp))
            if tj.get('size') == args.size and args.scaling == 'weak' and not args.generic:
                traffic = tj.get('dram_bytes_per_launch')
        # the bytes that really cross the HBM pins: ncu's DRAM counters when a capture of this
        # size exists, else the layout's own count (ncu confirms they agree to 1 %)
        actual_bytes = traffic if traffic is not None else layout_bytes * cells
        frac_actual = actual_bytes / (kernel_ms * 1e-3) / 1e9 / peak
        line = {
            'metric': 'grid cell-updates/sec (fp64 routing step)', 'value': value,
            'unit': 'cell-updates/s', 'n_gpus': D.world, 'steps': K, 'warmup': W,
            'ms_per_step': ms / K, 'higher_is_better': True, 'scaling': args.scaling,
            'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': workload_text(args, D.world),
                       'grid_per_gpu': [int(s.ny), int(s.nx)], 'wave': args.wave,
                       'noflow_fraction': noflow,
                       'pits': ('DEM not pit-filled (as FILL_PITS_IN_Z0 = 0 in the shipped cfgs): %d pits, each a '
                                'no-flow sink' % int(s.d8_stats['n_pits'])) if s.pits_filled is None else
                               ('DEM pit-filled by tf_fill_pits %s: %d pits left (grid edge)' % (
                                   json.dumps(s.pits_filled), int(s.d8_stats['n_pits']))),
                       'layout': 'generic-f64' if args.generic else 'packed (exact f32 params + angle table)',
                       'l2_policy': 'inputs larger than L2 (%.1f GB touched per step)' % (layout_bytes * cells / 1e9),
                       'parallelism': 'row strips x%d, halo rows over NCCL (%s)' % (
                           D.world, {'kinematic': 'Q once per step', 'diffusive': 'd between the phases, Q at the end',
                                     'dynamic': 'd between the phases, Q and u at the end'}[args.wave])},
            'parity': parity,
            'strong_scaling_dynamic': strong,
            'd8_area': None if area_s is None else {
                'seconds': area_s, 'cells_per_s': total_cells / area_s if area_s else None,
                'exchange_rounds': getattr(s, 'area_rounds', 0), 'unresolved_cells': int(s.area_unresolved),
                'global_cells': int(total_cells),
                'note': 'tf_d8_area, host call to host return; on strips: owner computes, edge rows of the area '
                        'exchanged until no flow path has anything left to hand over'},
            'e2e': {'value': e2e, 'unit': 'cell-updates/s', 'h2d_bytes_per_step': 8 * 6,
                    'd2h_bytes_per_step': 8 * 4 * s.n_outlets + 16},
            'e2e_grid_rain': None if e2e_grid is None else {
                'value': e2e_grid, 'unit': 'cell-updates/s', 'h2d_bytes_per_step': int(8 * cells),
                'd2h_bytes_per_step': 8 * 4 * s.n_outlets + 16 + 8,
                'note': 'a float64 rain grid uploaded from page-locked memory every step, R and its '
                        'integral formed on the device (tf_chan_set_forcing / tf_chan_assemble_R)'},
            'gpu_launches': int(launches),
            'clocks': clocks,
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                         'frac': achieved / peak, 'traffic': traffic, 'frac_actual': frac_actual,
                         'actual_bytes_source': 'ncu dram bytes (profiles/traffic_%s.json)' % args.wave
                                                if traffic is not None else 'layout bytes per cell x cells',
                         'peak_source': peak_src,
                         'algorithmic_bytes_per_cell': algo,
                         'layout_bytes_per_cell': layout_bytes,
                         'kernel_ms': kernel_ms},
            'wall_ms_timed_region': wall_ms,
        }
        if not args.no_cpu and D.world == 1:          # the CPU baseline is an N = 1 item
            v, el = cpu_path(args.wave, args.cpu_size, args.cpu_steps, 2)
            line['cpu_baseline'] = {'value': v, 'unit': 'cell-updates/s', 'cores': 1, 'kind': 'port',
                                    'sample': '%d steps of the same workload on a %dx%d grid (%.1f s), '
                                              'NumPy oracle, Python %d.%d; python2.7 %s' % (
                                                  args.cpu_steps, args.cpu_size, args.cpu_size, el,
                                                  sys.version_info[0], sys.version_info[1],
                                                  'found' if _have_py27() else 'unavailable on this box')}
        elif D.world > 1:
            line['cpu_baseline'] = None               # measured at N = 1 only
    s.close()
    D.close()
    if line is not None:
        print(json.dumps(line), file=_OUT, flush=True)


def _have_py27():
    try:
        return subprocess.run(['python2.7', '-c', 'import numpy'], capture_output=True, timeout=10).returncode == 0
    except Exception:
        return False


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    size = args.cpu_size
    # one "step" = one update of the reference algorithm on the bounded sample
    from oracle import channels as och  # noqa: F401  (the reference arm times the oracle port)
    v, el = cpu_path(args.wave, size, args.steps, max(args.warmup, 0))
    line = {
        'impl': 'reference', 'metric': 'grid cell-updates/sec (fp64 routing step)', 'value': v,
        'unit': 'cell-updates/s', 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': el / args.steps * 1e3, 'higher_is_better': True, 'scaling': args.scaling,
        'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': workload_text(args, args.gpus),
                   'grid_per_gpu': [args.size if args.scaling == 'weak' else args.size // max(args.gpus, 1),
                                    args.size], 'wave': args.wave,
                   'sample': 'each step = one update of the reference algorithm on a %dx%d sample of '
                             'that workload (same generator, same parameters), host cores only' % (size, size)},
        'cpu_baseline': {'value': v, 'unit': 'cell-updates/s', 'cores': 1, 'kind': 'port',
                         'sample': '%d steps on a %dx%d grid; NumPy restatement of channels_base.update() '
                                   '(the reference itself is Python 2.7 only; python2.7 %s)' % (
                                       args.steps, size, size,
                                       'found' if _have_py27() else 'unavailable on this box')},
        'e2e': {'value': v, 'unit': 'cell-updates/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line), file=_OUT, flush=True)


def _claim_stdout():
    """Point fd 1 at stderr for the whole run (NCCL and friends print banners to
    stdout) and return a file on the real stdout for the one JSON line."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), 'w')
    os.dup2(2, 1)
    return real


def main():
    global _OUT
    _OUT = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--size', type=int, default=16384, help='rows = columns per GPU')
    ap.add_argument('--wave', default='kinematic', choices=sorted(WAVE_ID))
    ap.add_argument('--generic', action='store_true', help='float64 parameter rasters (no packing)')
    ap.add_argument('--cpu-size', type=int, default=2048)
    ap.add_argument('--cpu-steps', type=int, default=25)
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-grid-rain', action='store_true', help='skip the extra e2e leg with a rain grid per step')
    ap.add_argument('--scaling', default='weak', choices=['weak', 'strong'],
                    help='weak: --size rows per GPU; strong: one --size x --size grid cut into N strips')
    ap.add_argument('--with-area', action='store_true', help='also compute (and time) the D8 contributing area')
    ap.add_argument('--fill-pits', action='store_true', help='one GPU: fill the depressions of the DEM first')
    ap.add_argument('--coupled', action='store_true',
                    help='configs[4]: met + snow + evap + infil providers on the device + kinematic routing')
    ap.add_argument('--no-parity', action='store_true', help='skip the strips / single strip / oracle check')
    ap.add_argument('--no-strong-leg', action='store_true',
                    help='N > 1: skip the extra strong-scaling line (dynamic wave, one grid cut into N strips)')
    ap.add_argument('--strong-size', type=int, default=16384)
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    elif args.coupled:
        run_coupled(args)
    else:
        run_gpu(args)


if __name__ == '__main__':
    main()
