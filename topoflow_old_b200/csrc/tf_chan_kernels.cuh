// tf_chan_kernels.cuh -- the routing kernels.
//
// Thread layout: a warp owns 64 consecutive cells of one row, TF_CPT = 2 per
// lane, so every raster access is a 128-bit (float64), 64-bit (float32) or
// 16-bit (uint8) load per lane and fully coalesced per warp.  A block is
// 32 x 8 threads = a tile of 64 columns x 8 rows, 4 blocks per SM at 64
// registers; vertical neighbours are mostly served by the block's own L1
// lines, the rest by L2.
//
// The eight-neighbour inflow uses the per-cell child mask built with the D8
// grid (tf_d8.cu:k_inmask): bit j set means "the neighbour opposite to
// direction j drains into me"; the adds are made in ascending j, which is the
// reference's k = 1..8 order (channels_base.py:1346-1361), each predicated on
// its bit.  The running sum stays in registers.
//
// Latency: a tile's loads form dependent chains (mask -> neighbour Q, angle
// index -> table, depth -> velocity factor).  Every block therefore asks L2
// (prefetch.global.L2, no registers held) for the cells of the tile one tile
// row further south, which a later wave of blocks will read: their loads are
// L2 hits instead of DRAM round trips (ChanArgs::pf_rows, TF_B200_PF_ROWS).
#pragma once
#include "tf_chan_math.cuh"

#ifndef TF_MINBLOCKS
#define TF_MINBLOCKS 4
#endif
#ifndef TF_MINBLOCKS_AB
#define TF_MINBLOCKS_AB 5
#endif
// tuning switches; the defaults are the measured best (profiles/sweep_r01.txt)
#ifndef TF_GATHER_UNROLL
#define TF_GATHER_UNROLL 1
#endif
#ifndef TF_LEAN_CHECKS
#define TF_LEAN_CHECKS 1
#endif
#ifndef TF_Q_SMEM
#define TF_Q_SMEM 0      /* 1: the tile's discharge (+ a one-cell ring) staged in shared memory for the inflow gather */
#endif
#define TF_SQW 68
#ifndef TF_GRID_CONSTANT
#define TF_GRID_CONSTANT 0
#endif
#if TF_GRID_CONSTANT
#define TF_GC __grid_constant__
#else
#define TF_GC
#endif
#ifndef TF_TILE_RCP
#define TF_TILE_RCP 1   /* depth divisions through precomputed reciprocals (exact, div_by_rcp) */
#endif
#ifndef TF_PREFETCH
#define TF_PREFETCH 9   /* bit 0: kfac of this tile, bit 3: look-ahead of a.pf_rows rows */
#endif

struct ChanArgs {
    int nx, ny, pitch;
    int y_begin, y_end;            // rows this launch covers
    int pf_rows;                   // L2 look-ahead distance in rows (0: off)
    int row_offset, global_ny;     // strip placement (bad-cell index, borders)
    double dt, rho;
    // state
    const double* Q;               // Q of this step (read with halo)
    double* Qn;                    // u * A_wet, next step's Q
    double* vol;
    double* d;
    double* u;                     // velocity written by this step
    const double* u_old;           // dynamic wave: last step's velocity
    long long steps_done;          // time steps taken since tf_chan_init
    const double* Rgrid;           // NULL -> Rs
    const double* Rrate;           // with Rgrid: R = Rrate * t_now + Rgrid (providers being interpolated)
    double t_now;                  // channel time at the start of this step [s]
    double Rs;
    const double* da_rows;         // [y], y = -1 .. ny
    // D8
    const uint8_t* code;
    const uint8_t* inmask;
    const float* dw;
    uint8_t* mstate;               // bit0: overbank rule has fired; bit1: fires next step
    int mut_width, mut_angle;
    // parameters, generic layout
    const double *w64, *tan64, *cos64, *S64, *n64, *ds64;
    // parameters, packed layout
    const float *w32, *S32, *n32, *ds32;
    const uint8_t* aidx;
    const double* lut;             // [256] tan, [256] cos, [256] 1 / (2 tan)
    const double* kfac;            // kinematic + Manning: sqrt(S_bed) / n, float64
    const float* len_rows;         // [16 * row] (see tf_handle_s::len_rows): ds {diag, E/W, N/S, -}, dw {diag, E/W, N/S, -}; NULL: read the rasters
    long long* bad;                // [0] depth, [1] velocity: min bad global index
    // marching kernels (tf_chan_march.cuh)
    double* vol_out;               // ping-pong partner of vol (== vol for the tile kernels)
    int march;                     // mstate holds two generations (see mstate_fired)
    int n_bands, seg_rows;         // column bands per row, rows per warp task
    int has_north, has_south;      // a neighbour strip owns row -1 / row ny
    int rcp_ok;                    // the reciprocal columns of len_rows / lut may be used
};

// Overbank state byte.  Tile kernels: bit 0 "the rule has fired", bit 1 "it
// fires at the next step".  The marching kernels keep two generations of that
// pair: generation g = steps_done & 1 lives in bits 2g and 2g + 1.  A step
// reads generation g and writes generation g ^ 1 (leaving g untouched), so a
// warp that recomputes a row owned by a neighbouring segment always finds the
// state that row had at the start of the step, whenever the owner writes.
__device__ __forceinline__ int mstate_fired(const ChanArgs& a, int byte) {
    return a.march ? ((byte >> (2 * (int)(a.steps_done & 1))) & 1) : (byte & 1);
}

// ---------------------------------------------------------------------------
template <bool PACKED>
__device__ __forceinline__ void load_shape(const ChanArgs& a, int64_t i, double (&w)[CPT],
                                           double (&t)[CPT], double (&c)[CPT], double (&ds)[CPT]) {
    if (PACKED) {
        ld4(a.w32 + i, w);
        if (!a.len_rows) ld4(a.ds32 + i, ds);
        int ai[CPT];
        ld4(a.aidx + i, ai);
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            t[k] = __ldg(a.lut + ai[k]);
            c[k] = __ldg(a.lut + 256 + ai[k]);
        }
    } else {
        ld4(a.w64 + i, w);
        ld4(a.ds64 + i, ds);
        ld4(a.tan64 + i, t);
        ld4(a.cos64 + i, c);
    }
}

// Flow length ds and flow width dw are functions of the D8 code and of three
// per-row lengths (update_flow_length_grid / update_flow_width_grid,
// d8_global.py:863-1004; ds then scaled by the sinuosity, channels_base.py:882).
// When the host has verified that for every cell (tf_chan_layout.cu), the
// packed layout reads the code it needs anyway and 32 bytes per row instead of
// two float32 rasters: 8 bytes per cell and step less from HBM.
template <bool PACKED>
__device__ __forceinline__ void load_lengths(const ChanArgs& a, int64_t i, int y, const int (&code)[CPT],
                                             double (&ds)[CPT], double (&dwv)[CPT]) {
    if (PACKED && a.len_rows) {
        const float4 rs = __ldg(reinterpret_cast<const float4*>(a.len_rows + 16 * y));
        const float4 rw = __ldg(reinterpret_cast<const float4*>(a.len_rows + 16 * y + 4));
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const int cd = code[k];                      // 1,4,16,64 diagonal; 2,32 E/W; 8,128 N/S; 0
            ds[k] = (double)((cd & 0x55) ? rs.x : ((cd & 0x88) ? rs.z : rs.y));
            dwv[k] = (double)((cd & 0x55) ? rw.x : ((cd & 0x22) ? rw.y : rw.z));
        }
    } else {
        ld4(a.dw + i, dwv);                              // ds came with load_shape
    }
}

template <bool PACKED>
__device__ __forceinline__ void load_slope_rough(const ChanArgs& a, int64_t i, double (&S)[CPT],
                                                 double (&n)[CPT]) {
    if (PACKED) {
        ld4(a.S32 + i, S);
        ld4(a.n32 + i, n);
    } else {
        ld4(a.S64 + i, S);
        ld4(a.n64 + i, n);
    }
}

__device__ __forceinline__ void prefetch_l2(const void* p) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// Software pipelining through L2: ask for the cells a later wave of blocks
// will read (pf_rows rows further south) so that its loads find them in L2
// instead of paying a DRAM round trip per dependent load.  No registers held.
template <bool PACKED>
__device__ __forceinline__ void prefetch_shape(const ChanArgs& a, int64_t j) {
    if (!(PACKED && a.len_rows)) prefetch_l2(a.dw + j);
    if (PACKED) {
        prefetch_l2(a.w32 + j);
        if (!a.len_rows) prefetch_l2(a.ds32 + j);
    } else {
        prefetch_l2(a.w64 + j); prefetch_l2(a.ds64 + j);
        prefetch_l2(a.tan64 + j); prefetch_l2(a.cos64 + j);
    }
}
template <bool PACKED>
__device__ __forceinline__ void prefetch_slope_rough(const ChanArgs& a, int64_t j) {
    if (PACKED) { prefetch_l2(a.S32 + j); prefetch_l2(a.n32 + j); }
    else { prefetch_l2(a.S64 + j); prefetch_l2(a.n64 + j); }
}
// what volume_and_depth reads (kinematic step, phase A)
template <bool PACKED>
__device__ __forceinline__ void prefetch_ahead_a(const ChanArgs& a, int64_t j) {
    prefetch_l2(a.Q + j);
    prefetch_l2(a.vol + j);
    prefetch_shape<PACKED>(a, j);
    if (a.Rgrid) prefetch_l2(a.Rgrid + j);
    if ((threadIdx.x & 7) == 0) {           // byte streams: one request per 32-byte sector
        prefetch_l2(a.code + j);
        prefetch_l2(a.inmask + j);
        prefetch_l2(a.mstate + j);
        if (PACKED) prefetch_l2(a.aidx + j);
    }
}

// overbank rule, channels_base.py:1449-1454: once "top_width > dw" has fired
// the cell is a rectangle as wide as the flow width
__device__ __forceinline__ void apply_overbank(const ChanArgs& a, int m, double dw, double& w,
                                               double& t, double& c) {
    if (m) {
        if (a.mut_width) w = dw;
        if (a.mut_angle) { t = 0.0; c = 1.0; }
    }
}

// offset of the child that bit j of the mask stands for
__device__ __forceinline__ int64_t child_offset(int j, int pitch) {
    // neighbour j is at (dr, dc); the child sits opposite: (-dr, -dc)
    const int dr = (j == 0 || j >= 6) ? -1 : ((j >= 2 && j <= 4) ? 1 : 0);
    const int dc = (j <= 2) ? 1 : ((j >= 4 && j <= 6) ? -1 : 0);
    return -((int64_t)dr * pitch + dc);
}

__device__ __forceinline__ void flag_bad(const ChanArgs& a, int which, int x, int y) {
    const long long gid = (long long)(y + a.row_offset) * a.nx + x;
    atomicMin(reinterpret_cast<long long*>(a.bad + which), gid);
}

// ---------------------------------------------------------------------------
// mass balance + trapezoid depth for CPT cells (update_flow_volume :1308-1375,
// update_flow_depth :1379-1532).  Returns the effective shape in w, t, c.
// ---------------------------------------------------------------------------
template <bool PACKED, bool RCP>
__device__ __forceinline__ void volume_and_depth(const ChanArgs& a, int x4, int y, int64_t i,
                                                 double (&Q)[CPT], double (&v)[CPT], double (&d)[CPT],
                                                 double (&w)[CPT], double (&t)[CPT], double (&c)[CPT],
                                                 double (&ds)[CPT], double (&dwv)[CPT], int (&code)[CPT],
                                                 int (&m)[CPT], int (&ms)[CPT], const double* sqc = nullptr) {
    int mask[CPT];
    ld4(a.Q + i, Q);
    ld4c(a.vol + i, v);
    ld4(a.code + i, code);
    ld4(a.inmask + i, mask);
    ld4c(a.mstate + i, ms);
    load_shape<PACKED>(a, i, w, t, c, ds);
    load_lengths<PACKED>(a, i, y, code, ds, dwv);
    double R[CPT];
    if (a.Rgrid) {
        ld4(a.Rgrid + i, R);
        if (a.Rrate) {                                     // a * t + b, time_interpolation.py:677
            double R1[CPT];
            ld4(a.Rrate + i, R1);
#pragma unroll
            for (int k = 0; k < CPT; ++k) R[k] = (R1[k] * a.t_now) + R[k];
        }
    } else {
#pragma unroll
        for (int k = 0; k < CPT; ++k) R[k] = a.Rs;
    }
    const double da = a.da_rows[y];
    const double dt = a.dt;
#if TF_TILE_RCP
    // (RCP: measured faster in phase A, slower -- a spill -- in the kinematic kernel)
    int aidx_of[CPT];
#pragma unroll
    for (int k = 0; k < CPT; ++k) aidx_of[k] = 0;
    if (RCP && PACKED && a.len_rows && a.rcp_ok) ld4(a.aidx + i, aidx_of);
#endif
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        m[k] = (ms[k] & 1) | ((ms[k] >> 1) & 1);
        apply_overbank(a, m[k], dwv[k], w[k], t[k], c[k]);
        double acc = v[k] + (R[k] * da) * dt;                  // :1334
        int mk = mask[k];
#if TF_Q_SMEM
        {
            // the same eight adds, the children's discharge read from the block's shared tile
            const double* q = sqc + k;
            if (mk & 1) acc += dt * q[TF_SQW - 1];
            if (mk & 2) acc += dt * q[-1];
            if (mk & 4) acc += dt * q[-TF_SQW - 1];
            if (mk & 8) acc += dt * q[-TF_SQW];
            if (mk & 16) acc += dt * q[-TF_SQW + 1];
            if (mk & 32) acc += dt * q[1];
            if (mk & 64) acc += dt * q[TF_SQW + 1];
            if (mk & 128) acc += dt * q[TF_SQW];
        }
#elif TF_GATHER_UNROLL
        // the k = 1..8 adds of :1346-1361 in order, each predicated on its bit
        // of the mask; offsets are compile-time forms of the pitch, and the
        // east / west child of a cell pair is already in registers
        {
            const double* q = a.Q + i + k;
            const int64_t p = a.pitch;
            if (mk & 1) acc += dt * __ldg(q + (p - 1));
            if (mk & 2) acc += dt * ((CPT == 2 && k == 1) ? Q[0] : __ldg(q - 1));
            if (mk & 4) acc += dt * __ldg(q - (p + 1));
            if (mk & 8) acc += dt * __ldg(q - p);
            if (mk & 16) acc += dt * __ldg(q - (p - 1));
            if (mk & 32) acc += dt * ((CPT == 2 && k == 0) ? Q[1] : __ldg(q + 1));
            if (mk & 64) acc += dt * __ldg(q + (p + 1));
            if (mk & 128) acc += dt * __ldg(q + p);
        }
#else
        while (mk) {                                           // :1346-1361
            const int j = __ffs(mk) - 1;
            mk &= mk - 1;
            acc += dt * __ldg(a.Q + i + k + child_offset(j, a.pitch));
        }
#endif
        acc -= Q[k] * dt;                                      // :1366
        acc = np_max0(acc);                                    // :1375
        v[k] = acc;
#if TF_TILE_RCP
        double dd;
        if (RCP && PACKED && a.len_rows && a.rcp_ok) {
            // same bits as chan_depth (div_by_rcp is exact): the reciprocals come
            // from the row table and the angle table instead of two divisions
            const double* rr = reinterpret_cast<const double*>(a.len_rows + 16 * y + 8);
            const int cd = code[k];
            const double rcp_ds = __ldg(rr + ((cd & 0x55) ? 0 : ((cd & 0x88) ? 2 : 1)));
            const double rcp_den = m[k] && a.mut_angle ? 0.0 : __ldg(a.lut + 512 + aidx_of[k]);
            dd = chan_depth_rcp(acc, w[k], t[k], ds[k], rcp_ds, rcp_den);
        } else dd = chan_depth(acc, w[k], t[k], ds[k]);
#else
        double dd = chan_depth(acc, w[k], t[k], ds[k]);
#endif
        if (code[k] == 0) dd = 0.0;                            // :1511
        d[k] = np_max0(dd);                                    // :1526
    }
    (void)x4;
}

// next step's overbank test with this step's depth (:1449-1450), packed into
// the state byte; written only when it changes
__device__ __forceinline__ void store_mstate(const ChanArgs& a, int64_t i, const int (&ms)[CPT],
                                             const int (&m)[CPT], const double (&d)[CPT],
                                             const double (&w)[CPT], const double (&t)[CPT],
                                             const double (&dwv)[CPT]) {
    int nw[CPT];
    bool changed = false;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const double top = w[k] + ((2.0 * d[k]) * t[k]);
        nw[k] = m[k] | ((top > dwv[k]) ? 2 : 0);
        changed = changed || (nw[k] != ms[k]);
    }
    if (changed) st4(a.mstate + i, nw);
}

#if TF_Q_SMEM
// Stage rows y0-1 .. y0+8, columns x0-1 .. x0+64 of Q in shared memory; returns the pointer to
// this thread's first cell in the tile.  Every thread of the block must call it (barrier).
__device__ __forceinline__ const double* stage_Q_tile(const ChanArgs& a, double (*sq)[TF_SQW], int x4, int y) {
    const int r = threadIdx.y, lane = threadIdx.x;
    const double* Q = a.Q;
    if (y <= a.ny) {                                        // (row ny is an apron row: valid memory)
        const int64_t i = (int64_t)y * a.pitch + x4;
        const double2 q = *reinterpret_cast<const double2*>(Q + i);
        sq[r + 1][1 + 2 * lane] = q.x; sq[r + 1][2 + 2 * lane] = q.y;
        if (lane == 0) sq[r + 1][0] = Q[i - 1];
        if (lane == 31) sq[r + 1][65] = Q[i + 2];
        if (r == 0) {
            const double2 n = *reinterpret_cast<const double2*>(Q + i - a.pitch);
            sq[0][1 + 2 * lane] = n.x; sq[0][2 + 2 * lane] = n.y;
            if (lane == 0) sq[0][0] = Q[i - a.pitch - 1];
            if (lane == 31) sq[0][65] = Q[i - a.pitch + 2];
        }
        if (r == 7 || y == a.ny) {
            const double2 n = *reinterpret_cast<const double2*>(Q + i + a.pitch);
            sq[r + 2][1 + 2 * lane] = n.x; sq[r + 2][2 + 2 * lane] = n.y;
            if (lane == 0) sq[r + 2][0] = Q[i + a.pitch - 1];
            if (lane == 31) sq[r + 2][65] = Q[i + a.pitch + 2];
        }
    }
    __syncthreads();
    return &sq[r + 1][1 + 2 * lane];
}
#endif

// ---------------------------------------------------------------------------
// kinematic wave: the whole step in one pass
// ---------------------------------------------------------------------------
// KF: S_bed and n never change in the kinematic wave, so Manning's
// (Rh^(2/3) * sqrt(S)) / n is evaluated as Rh^(2/3) * (sqrt(S)/n) with the
// factor precomputed once on the host (one float64 read instead of two
// float32 reads, a square root and a division per cell and step).  The
// regrouping moves u by at most 1 ulp, next to the ~2 ulp of the cube root.
template <bool PACKED, int LAW, bool KF>
__global__ void __launch_bounds__(256, TF_MINBLOCKS)
k_step_kinematic(const TF_GC ChanArgs a) {
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * CPT;
    const int y = a.y_begin + blockIdx.y * 8 + threadIdx.y;
#if TF_Q_SMEM
    __shared__ double sq[10][TF_SQW];
    const double* sqc = stage_Q_tile(a, sq, x4, y);
#else
    const double* sqc = nullptr;
#endif
    if (x4 >= a.pitch || y >= a.y_end) return;
    const int64_t i = (int64_t)y * a.pitch + x4;
    double Q[CPT], v[CPT], d[CPT], w[CPT], t[CPT], c[CPT], ds[CPT], dwv[CPT], S[CPT], n[CPT], u[CPT], Qn[CPT];
    int code[CPT], m[CPT], ms[CPT];
#if TF_PREFETCH & 1
    // the velocity factor is read after the depth arithmetic (no registers to
    // hold it until then): start its trip from DRAM now, so that read hits L2
    if (KF) prefetch_l2(a.kfac + i);
#endif
#if TF_PREFETCH & 2
    // the tile's halo rows of Q, which no warp of this block loads as its own
    if (threadIdx.y == 0) prefetch_l2(a.Q + i - a.pitch);
    if (threadIdx.y == 7) prefetch_l2(a.Q + i + a.pitch);
#endif
#if TF_PREFETCH & 8
    if (a.pf_rows > 0 && y + a.pf_rows < a.y_end) {
        const int64_t j = i + (int64_t)a.pf_rows * a.pitch;
        prefetch_ahead_a<PACKED>(a, j);
        if (KF) prefetch_l2(a.kfac + j);
        else prefetch_slope_rough<PACKED>(a, j);
    }
#endif
    volume_and_depth<PACKED, false>(a, x4, y, i, Q, v, d, w, t, c, ds, dwv, code, m, ms, sqc);
    if (KF) ld4(a.kfac + i, S);
    else load_slope_rough<PACKED>(a, i, S, n);
    bool bad_d = false, bad_u = false;
    int bx = 0;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const bool noflow = (code[k] == 0);
        const XSection xs = chan_xsection_step(d[k], w[k], t[k], c[k], noflow);
        double uu = KF ? pow_two_thirds(xs.Rh) * S[k]
                       : chan_velocity<LAW>(xs.Rh, S[k], n[k], d[k]);   // kinematic: S = S_bed
        if (noflow) uu = 0.0;                                      // :1783-1791
        u[k] = uu;
        Qn[k] = uu * xs.A_wet;                                     // :1188 of the next step
#if TF_LEAN_CHECKS
        // d = np_max0(.) is never negative, so "bad" is NaN or +inf; padding
        // cells (x >= nx) hold a harmless shape with code 0 and stay at d = u = 0
        bad_d = bad_d || !(d[k] <= 1.7976931348623157e308);
        bad_u = bad_u || !(uu >= 0.0 && uu <= 1.7976931348623157e308);
#else
        if (x4 + k < a.nx) {
            if (bad_value(d[k]) && !bad_d) { bad_d = true; bx = x4 + k; }
            if (bad_value(uu) && !bad_u) { bad_u = true; if (!bad_d) bx = x4 + k; }
        }
#endif
    }
    st4(a.vol + i, v);
    st4(a.d + i, d);
    st4(a.u + i, u);
    st4(a.Qn + i, Qn);
    store_mstate(a, i, ms, m, d, w, t, dwv);
    if (bad_d || bad_u) {
        // report the first offending cell of each kind in this segment
#pragma unroll
        for (int k = CPT - 1; k >= 0; --k) {
            if (x4 + k < a.nx) {
                if (bad_value(d[k])) flag_bad(a, 0, x4 + k, y);
                if (bad_value(u[k])) flag_bad(a, 1, x4 + k, y);
            }
        }
    }
    (void)bx;
}

// ---------------------------------------------------------------------------
// diffusive / dynamic wave, phase A: volume and depth of every cell
// ---------------------------------------------------------------------------
template <bool PACKED>
__global__ void __launch_bounds__(256, TF_MINBLOCKS_AB)
k_phase_a(const TF_GC ChanArgs a) {
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * CPT;
    const int y = a.y_begin + blockIdx.y * 8 + threadIdx.y;
#if TF_Q_SMEM
    __shared__ double sq[10][TF_SQW];
    const double* sqc = stage_Q_tile(a, sq, x4, y);
#else
    const double* sqc = nullptr;
#endif
    if (x4 >= a.pitch || y >= a.y_end) return;
    const int64_t i = (int64_t)y * a.pitch + x4;
    double Q[CPT], v[CPT], d[CPT], w[CPT], t[CPT], c[CPT], ds[CPT], dwv[CPT];
    int code[CPT], m[CPT], ms[CPT];
#if TF_PREFETCH & 8
    if (a.pf_rows > 0 && y + a.pf_rows < a.y_end)
        prefetch_ahead_a<PACKED>(a, i + (int64_t)a.pf_rows * a.pitch);
#endif
    volume_and_depth<PACKED, true>(a, x4, y, i, Q, v, d, w, t, c, ds, dwv, code, m, ms, sqc);
    st4(a.vol + i, v);
    st4(a.d + i, d);
    store_mstate(a, i, ms, m, d, w, t, dwv);
#pragma unroll
    for (int k = 0; k < CPT; ++k)
        if (x4 + k < a.nx && bad_value(d[k])) flag_bad(a, 0, x4 + k, y);
}
