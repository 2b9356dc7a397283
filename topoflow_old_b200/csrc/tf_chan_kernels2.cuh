// tf_chan_kernels2.cuh -- phase B of the diffusive / dynamic wave and the
// small kernels around the step (outlet taps, diversions, derived fields,
// reductions).
#pragma once
#include "tf_chan_kernels.cuh"

// offset of the parent (downstream cell) of a cell with D8 code `code`
__device__ __forceinline__ int64_t parent_offset(int code, int pitch) {
    const int j = __ffs(code) - 1;
    return -child_offset(j, pitch);
}

// free-surface slope, channels_base.py:1536-1543.  A no-flow cell's parent is
// cell (0, 0) (d8_global.py:641-644), itself a no-flow cell with d = 0.
__device__ __forceinline__ double free_slope(const ChanArgs& a, int64_t idx, int code, double d,
                                             double S_bed, double ds) {
    const double dp = code ? a.d[idx + parent_offset(code, a.pitch)] : 0.0;
    return S_bed + ((d - dp) / ds);
}

// ---------------------------------------------------------------------------
// phase B: slope, velocity and next discharge once every depth is known
//   WAVE 1: channels_diffusive_wave.py:83-130
//   WAVE 2: channels_dynamic_wave.py:92-341
// ---------------------------------------------------------------------------
// (the dynamic wave's momentum balance holds more values at once: 4 blocks per
// SM at 64 registers without spills measured 3 % faster than 5 at 48 with)
template <bool PACKED, int LAW, int WAVE>
__global__ void __launch_bounds__(256, (WAVE == 2) ? TF_MINBLOCKS_AB - 1 : TF_MINBLOCKS_AB)
k_phase_b(const TF_GC ChanArgs a) {
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * CPT;
    const int y = a.y_begin + blockIdx.y * 8 + threadIdx.y;
    if (x4 >= a.pitch || y >= a.y_end) return;
    const int64_t i = (int64_t)y * a.pitch + x4;
    double d[CPT], w[CPT], t[CPT], c[CPT], ds[CPT], dwv[CPT], S[CPT], n[CPT], u[CPT], Qn[CPT];
    int code[CPT], ms[CPT];
#if TF_PREFETCH & 8
    if (a.pf_rows > 0 && y + a.pf_rows < a.y_end) {
        const int64_t j = i + (int64_t)a.pf_rows * a.pitch;
        prefetch_l2(a.d + j);
        prefetch_shape<PACKED>(a, j);
        prefetch_slope_rough<PACKED>(a, j);
        if (WAVE == 2) { prefetch_l2(a.Q + j); prefetch_l2(a.vol + j); prefetch_l2(a.u_old + j); }
        if ((threadIdx.x & 7) == 0) {
            prefetch_l2(a.code + j);
            prefetch_l2(a.mstate + j);
            if (PACKED) prefetch_l2(a.aidx + j);
            if (WAVE == 2) prefetch_l2(a.inmask + j);
        }
    }
#endif
    ld4c(a.d + i, d);
    ld4(a.code + i, code);
    ld4c(a.mstate + i, ms);
    load_shape<PACKED>(a, i, w, t, c, ds);
    load_lengths<PACKED>(a, i, y, code, ds, dwv);
    load_slope_rough<PACKED>(a, i, S, n);
    double Q[CPT], v[CPT], uo[CPT];
    int mask[CPT];
    if (WAVE == 2) {
        ld4(a.Q + i, Q);
        ld4c(a.vol + i, v);
        ld4(a.u_old + i, uo);
        ld4(a.inmask + i, mask);
    }
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const bool noflow = (code[k] == 0);
        apply_overbank(a, ms[k] & 1, dwv[k], w[k], t[k], c[k]);
        // the dynamic wave needs the wetted perimeter itself (:166)
        const XSection xs = (WAVE == 1) ? chan_xsection_step(d[k], w[k], t[k], c[k], noflow)
                                        : chan_xsection(d[k], w[k], t[k], c[k], noflow);
        const double dpar = code[k] ? a.d[i + k + parent_offset(code[k], a.pitch)] : 0.0;
        double s_free = free_slope_step(S[k], d[k], dpar, ds[k], WAVE == 2);
        double uu;
        if (WAVE == 1) {
            s_free = np_max0(s_free);                              // diffusive :118
            uu = chan_velocity<LAW>(xs.Rh, s_free, n[k], d[k]);
        } else {
            // momentum balance, channels_dynamic_wave.py:126-341
            const double dt = a.dt;
            const double u0 = (d[k] > 0.0) ? uo[k] : 0.0;          // :126-127
            const double vinv = (v[k] > 0.0) ? (1.0 / v[k]) : 0.0; // :140-145
            const double Ap = ds[k] * xs.P_wet;                    // :166
            const double Qtrm = u0 * Q[k];                         // :171
            const double grav = (TF_G * s_free) * v[k];            // :178
            const double f = chan_friction<LAW>(d[k], n[k]);
            const double fric = ((f * u0) * u0) * Ap;              // :190
            const double dM = ((grav - fric) - Qtrm) * dt;         // :214
            double M = (u0 * v[k]) + dM;                           // :222
            int mk = mask[k];
            // (a loop: with three loads per child it is faster than eight predicated terms)
            while (mk) {                                           // :244-259
                const int j = __ffs(mk) - 1;
                mk &= mk - 1;
                const int64_t ci = i + k + child_offset(j, a.pitch);
                const double uc = (a.d[ci] > 0.0) ? __ldg(a.u_old + ci) : 0.0;
                M += (uc * __ldg(a.Q + ci)) * dt;
            }
            M = np_max0(M);                                        // :312
            uu = np_max0(M * vinv);                                // :318-331
        }
        if (noflow) uu = 0.0;                                      // channels_base.py:1783-1791
        u[k] = uu;
        Qn[k] = uu * xs.A_wet;
        if (x4 + k < a.nx && bad_value(uu)) flag_bad(a, 1, x4 + k, y);
    }
    st4(a.u + i, u);
    st4(a.Qn + i, Qn);
}

// ---------------------------------------------------------------------------
// outlet taps: Q, u, d, f at the monitored cells (update_outlet_values,
// channels_base.py:1812-1832); canals_in_Q is written by k_diversions.
// ---------------------------------------------------------------------------
template <bool PACKED, int LAW>
__global__ void k_taps(const ChanArgs a, const int64_t* __restrict__ idx, int n, double* out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int64_t i = idx[k];
    const double d = a.d[i];
    const double rough = PACKED ? (double)a.n32[i] : a.n64[i];
    out[4 * k + 0] = a.Q[i];
    out[4 * k + 1] = a.u[i];
    out[4 * k + 2] = d;
    out[4 * k + 3] = chan_friction<LAW>(d, rough);
}

// ---------------------------------------------------------------------------
// diversions (update_diversions, channels_base.py:1226-1297), one thread:
// NumPy's "a[ids] += v" reads every a[ids] first and then assigns in order,
// so a repeated id keeps only its last update.  tmp holds the staged values.
// ---------------------------------------------------------------------------
struct PointsArgs {
    int n[4];
    const int64_t* idx[4];
    const double* val[4];
    double* tmp;            // 2 * max(n) staged values
    double* canals_in_out;  // [n canals] or NULL
};

__device__ __forceinline__ void fancy_add(double* a, const int64_t* idx, const double* val, int n,
                                          double sign, double scale, double* tmp) {
    for (int k = 0; k < n; ++k) tmp[k] = a[idx[k]] + sign * (val[k] * scale);
    for (int k = 0; k < n; ++k) a[idx[k]] = tmp[k];
}

static __global__ void k_diversions(PointsArgs p, double* Q, double* vol, double dt) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    double* tmp = p.tmp;
    if (p.n[TF_PT_SOURCES] > 0) {                                  // :1264-1266
        fancy_add(Q, p.idx[0], p.val[0], p.n[0], 1.0, 1.0, tmp);
        fancy_add(vol, p.idx[0], p.val[0], p.n[0], 1.0, dt, tmp);
    }
    if (p.n[TF_PT_SINKS] > 0) {                                    // :1271-1273
        fancy_add(Q, p.idx[1], p.val[1], p.n[1], -1.0, 1.0, tmp);
        fancy_add(vol, p.idx[1], p.val[1], p.n[1], -1.0, dt, tmp);
    }
    const int nc = p.n[TF_PT_CANALS_IN];
    if (nc > 0) {                                                  // :1278-1297
        double* cin = tmp + nc;
        for (int k = 0; k < nc; ++k) {
            cin[k] = p.val[2][k] * Q[p.idx[2][k]];                 // :1283
            if (p.canals_in_out) p.canals_in_out[k] = cin[k];
        }
        fancy_add(Q, p.idx[2], cin, nc, -1.0, 1.0, tmp);           // :1289-1290
        fancy_add(vol, p.idx[2], cin, nc, -1.0, dt, tmp);
        fancy_add(Q, p.idx[3], p.val[3], p.n[3], 1.0, 1.0, tmp);   // :1296-1297
        fancy_add(vol, p.idx[3], p.val[3], p.n[3], 1.0, dt, tmp);
    }
}

// ---------------------------------------------------------------------------
// first overbank test, from the initial depth (the reference evaluates it at
// the top of the first update_flow_depth with d = d0)
// ---------------------------------------------------------------------------
template <bool PACKED>
__global__ void __launch_bounds__(256)
k_init_mstate(const ChanArgs a) {
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * CPT;
    const int y = a.y_begin + blockIdx.y * 8 + threadIdx.y;
    if (x4 >= a.pitch || y >= a.y_end) return;
    const int64_t i = (int64_t)y * a.pitch + x4;
    double d[CPT], w[CPT], t[CPT], c[CPT], ds[CPT], dwv[CPT];
    ld4c(a.d + i, d);
    load_shape<PACKED>(a, i, w, t, c, ds);
    ld4(a.dw + i, dwv);
    // A depth set in the middle of a run (tf_chan_set_state after steps were
    // taken) re-evaluates the pending test only: a cell the rule has already
    // rewritten stays rewritten (channels_base.py:1451-1454 is permanent) and
    // is tested with its rewritten shape.
    int old[CPT];
#pragma unroll
    for (int k = 0; k < CPT; ++k) old[k] = 0;
    if (a.steps_done > 0) ld4c(a.mstate + i, old);
    int nw[CPT];
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const int fired = mstate_fired(a, old[k]);
        apply_overbank(a, fired, dwv[k], w[k], t[k], c[k]);
        const double top = w[k] + ((2.0 * d[k]) * t[k]);
        nw[k] = fired | ((top > dwv[k]) ? 2 : 0);
        if (a.march) nw[k] |= nw[k] << 2;        // both generations start equal
    }
    st4(a.mstate + i, nw);
}

// ---------------------------------------------------------------------------
// derived fields on demand (tf_chan_get)
// ---------------------------------------------------------------------------
template <bool PACKED, int LAW>
__global__ void __launch_bounds__(256)
k_derive(const ChanArgs a, int which, int wave, double* out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= a.nx || y >= a.ny) return;
    const int64_t i = (int64_t)y * a.pitch + x;
    const int code = a.code[i];
    const bool noflow = (code == 0);
    const int m = mstate_fired(a, a.mstate[i]);
    double w, t, c, ds, S, n;
    if (PACKED) {
        w = a.w32[i]; ds = a.ds32[i]; S = a.S32[i]; n = a.n32[i];
        t = a.lut[a.aidx[i]]; c = a.lut[256 + a.aidx[i]];
    } else {
        w = a.w64[i]; ds = a.ds64[i]; S = a.S64[i]; n = a.n64[i];
        t = a.tan64[i]; c = a.cos64[i];
    }
    apply_overbank(a, m, (double)a.dw[i], w, t, c);
    const double d = a.d[i];
    const XSection xs = chan_xsection(d, w, t, c, noflow);
    double r = 0.0;
    switch (which) {
        case TF_F_F: r = chan_friction<LAW>(d, n); break;
        case TF_F_RH: r = xs.Rh; break;
        case TF_F_A_WET: r = xs.A_wet; break;
        case TF_F_P_WET: r = xs.P_wet; break;
        case TF_F_TAU:
        case TF_F_U_STAR: {
            // update_shear_stress :1567-1577 (|S_free| before the diffusive clip)
            const double slope = (wave == 0) ? S : fabs(free_slope(a, i, code, d, S, ds));
            const double tau = ((a.rho * TF_G) * d) * slope;
            r = (which == TF_F_TAU) ? tau : sqrt(tau / a.rho);     // :1586
        } break;
        case TF_F_FROUDE: r = (d > 0.0) ? a.u[i] / sqrt(TF_G * d) : 0.0; break;   // :1807
        case TF_F_WIDTH: r = w; break;
        case TF_F_MUTATED: r = (double)m; break;
        case TF_F_S_FREE: {
            // not stored: S_bed + (d - d[parent]) / ds (:1542-1543), clipped at 0 by the
            // diffusive wave (channels_diffusive_wave.py:118); S_bed before the first step (:889)
            const double sf = free_slope(a, i, code, d, S, ds);
            r = (wave == 0 || a.steps_done == 0) ? S : ((wave == 1) ? np_max0(sf) : sf);
        } break;
        default: break;
    }
    out[i] = r;
}

// ---------------------------------------------------------------------------
// reductions: per-block partials, summed / compared on the host in block
// order (deterministic)
// ---------------------------------------------------------------------------
static __global__ void __launch_bounds__(256)
k_sum_R(const ChanArgs a, double* partial) {
    // sum of (R*da)*dt over one row segment per block: update_R_integral,
    // channels_base.py:1151-1160
    __shared__ double sh[256];
    const int y = blockIdx.x;
    double acc = 0.0;
    const double da = a.da_rows[y];
    for (int x = threadIdx.x; x < a.nx; x += blockDim.x)
        acc += (a.Rgrid[(int64_t)y * a.pitch + x] * da) * a.dt;
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[y] = sh[0];
}

struct ForcingArgs {
    const double* grid[TF_R_COUNT];   // cell (0,0) of each flux raster, or NULL
    double scalar[TF_R_COUNT];
};

// update_R (:1103-1147) and, in the same pass, the row sums of (R*da)*dt for
// update_R_integral (:1151-1160).  One block per row.
static __global__ void __launch_bounds__(256)
k_assemble_R(const ChanArgs a, const ForcingArgs fa, double* __restrict__ R, double* partial) {
    __shared__ double sh[256];
    const int y = blockIdx.x;
    const double da = a.da_rows[y];
    double acc = 0.0;
    for (int x = threadIdx.x; x < a.nx; x += blockDim.x) {
        const int64_t i = (int64_t)y * a.pitch + x;
        double v[TF_R_COUNT];
#pragma unroll
        for (int k = 0; k < TF_R_COUNT; ++k) v[k] = fa.grid[k] ? __ldcs(fa.grid[k] + i) : fa.scalar[k];
        const double r = (((v[TF_R_P_RAIN] + v[TF_R_SM]) + v[TF_R_GW]) + v[TF_R_MR])
                         - (v[TF_R_ET] + v[TF_R_IN]);
        R[i] = r;
        acc += (r * da) * a.dt;
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[y] = sh[0];
}

static __global__ void __launch_bounds__(256)
k_minmax(const ChanArgs a, double* partial) {
    // interior min / max of Q, u, d (update_mins_and_maxes :1907-1916):
    // one block per interior row, partial[6 * row + ...]
    __shared__ double sh[6][256];
    const int y = blockIdx.x;
    const int gy = y + a.row_offset;
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    bool nan_seen[3] = {false, false, false};
    if (gy >= 1 && gy <= a.global_ny - 2) {
        for (int x = 1 + threadIdx.x; x <= a.nx - 2; x += blockDim.x) {
            const int64_t i = (int64_t)y * a.pitch + x;
            const double val[3] = {a.Q[i], a.u[i], a.d[i]};
#pragma unroll
            for (int f = 0; f < 3; ++f) {
                if (val[f] != val[f]) nan_seen[f] = true;
                lo[f] = fmin(lo[f], val[f]);
                hi[f] = fmax(hi[f], val[f]);
            }
        }
    }
#pragma unroll
    for (int f = 0; f < 3; ++f) {
        sh[2 * f][threadIdx.x] = nan_seen[f] ? NAN : lo[f];
        sh[2 * f + 1][threadIdx.x] = nan_seen[f] ? NAN : hi[f];
    }
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
#pragma unroll
            for (int f = 0; f < 6; ++f) {
                const double p = sh[f][threadIdx.x], q = sh[f][threadIdx.x + s];
                double r;
                if (p != p || q != q) r = NAN;
                else r = (f & 1) ? (p > q ? p : q) : (p < q ? p : q);
                sh[f][threadIdx.x] = r;
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0)
        for (int f = 0; f < 6; ++f) partial[6 * y + f] = sh[f][0];
}
