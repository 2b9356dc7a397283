// tf_chan_math.cuh -- per-cell arithmetic of the channel step.
//
// Every expression keeps the association NumPy used in the reference
// (left to right, no fused multiply-add: this TU is built with -fmad=false),
// so vol and d are reproduced bit for bit from bit-identical inputs.  The
// only operations that are not correctly rounded are the cube root (the
// reference's x**(1/3), x**(2/3)) and log, both within ~2 ulp.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/topoflow_b200.h"

#define TF_G 9.81                    // channels_base.py:456
#define TF_AVAL 0.476                // :457
#define TF_KAPPA 0.408               // :458

// np.maximum(x, 0): NaN propagates (fmax would swallow it)
__device__ __forceinline__ double np_max0(double x) {
    return (x != x) ? x : (x > 0.0 ? x : 0.0);
}
__device__ __forceinline__ double np_maxv(double x, double lo) {
    return (x != x) ? x : (x > lo ? x : lo);
}

// update_flow_depth, channels_base.py:1465-1486
__device__ __forceinline__ double chan_depth(double vol, double w, double t, double ds) {
    if (t == 0.0) {
        return vol / (w * ds);
    }
    const double denom = 2.0 * t;
    double arg = ((2.0 * denom) * vol) / ds;
    arg += w * w;
    return (sqrt(arg) - w) / denom;
}

// a / b, correctly rounded, from y = RN(1 / b) computed once (Markstein): with
// q0 = RN(a y) and the exact residual r = a - b q0 (one fused multiply-add),
// RN(q0 + r y) IS the IEEE quotient -- bit for bit what "a / b" returns, at 3
// instructions instead of ~16.  (Checked against a / b on 6e8 random pairs on
// the host, including all-ones significands.)  The residual is exact only
// while nothing underflows or overflows, so operands outside 2^-767 .. 2^768
// (and infinities / NaN) take the ordinary division; zero stays on the fast
// path (0 y = 0, r = 0).
static __device__ __noinline__ double div_ieee(double a, double b) { return a / b; }   // rare: kept out of line
__device__ __forceinline__ double div_by_rcp(double a, double b, double y) {
    const double q0 = a * y;
    const double r = __fma_rn(-q0, b, a);
    double q1 = __fma_rn(r, y, q0);
    const unsigned hi = (unsigned)__double2hiint(a) & 0x7fffffffu;
    const bool fast = ((hi - 0x10000000u) < 0x60000000u) || (a == 0.0);
    if (!fast) q1 = div_ieee(a, b);
    return q1;
}

// chan_depth with the reciprocals of ds and of 2 tan(angle) at hand (the
// marching kernels keep them per row / per table entry): same bits.
__device__ __forceinline__ double chan_depth_rcp(double vol, double w, double t, double ds,
                                                 double rcp_ds, double rcp_denom) {
    if (t == 0.0) {
        return vol / (w * ds);
    }
    const double denom = 2.0 * t;
    double arg = div_by_rcp((2.0 * denom) * vol, ds, rcp_ds);
    arg += w * w;
    return div_by_rcp(sqrt(arg) - w, denom, rcp_denom);
}

// x / y to within an ulp or two for the velocity path (u is a few ulp from
// the reference's pow in any case): float reciprocal as a seed, two Newton
// steps in float64.  y must be positive and inside the float range.
__device__ __forceinline__ double div_approx(double x, double y) {
    const double r0 = (double)__frcp_rn((float)y);
    const double r1 = __fma_rn(r0, __fma_rn(-y, r0, 1.0), r0);
    const double r2 = __fma_rn(r1, __fma_rn(-y, r1, 1.0), r1);
    double q = x * r2;
    if (!(y > 1e-30 && y < 1e30)) q = div_ieee(x, y);
    return q;
}

// update_trapezoid_Rh, channels_base.py:1615-1643
struct XSection { double A_wet, P_wet, Rh; };
__device__ __forceinline__ XSection chan_xsection(double d, double w, double t, double c, bool noflow) {
    XSection s;
    const double L2 = d * t;
    s.A_wet = d * (w + L2);
    s.P_wet = (c == 1.0) ? (w + (2.0 * d)) : (w + ((2.0 * d) / c));   // x / 1 is exact
    if (noflow) s.P_wet = 1.0;
    s.Rh = s.A_wet / s.P_wet;
    if (noflow) s.Rh = 0.0;
    return s;
}

// Wetted area and hydraulic radius for the velocity of the step kernels.
// A_wet is the reference's expression (it becomes Q, hence vol: exact).  Rh
// only feeds u, which is a few ulp from the reference's pow in any case:
// A / (w + 2d/c) is evaluated as (A c) / (w c + 2d) -- one division, not two.
#ifndef TF_RH_ONE_DIV
#define TF_RH_ONE_DIV 1
#endif
#ifndef TF_RH_APPROX_DIV
#define TF_RH_APPROX_DIV 1
#endif
#ifndef TF_FAST_MANNING
#define TF_FAST_MANNING 1   /* Manning's division by n and the free-surface slope's by ds through div_approx */
#endif
#ifndef TF_FAST_SEED
#define TF_FAST_SEED 1      /* cube-root seed from lg2.approx / ex2.approx instead of rcbrtf */
#endif
__device__ __forceinline__ XSection chan_xsection_step(double d, double w, double t, double c, bool noflow) {
#if TF_RH_ONE_DIV
    XSection s;
    const double L2 = d * t;
    s.A_wet = d * (w + L2);
    s.P_wet = 0.0;     // not used by the callers
#if TF_RH_APPROX_DIV
    s.Rh = noflow ? 0.0 : div_approx(s.A_wet * c, __fma_rn(w, c, 2.0 * d));
#else
    s.Rh = noflow ? 0.0 : ((s.A_wet * c) / __fma_rn(w, c, 2.0 * d));
#endif
    return s;
#else
    return chan_xsection(d, w, t, c, noflow);
#endif
}

// x ** (2/3) for x >= 0 (NaN and inf pass through like pow)
#ifndef TF_FAST_P23
#define TF_FAST_P23 1
#endif
// x^(-1/3) to ~1e-6 relative: enough for the two-term correction below (the
// first neglected term is 14/81 h^3 with h ~ 3e-6)
__device__ __forceinline__ double rcbrt_seed(double x) {
#if TF_FAST_SEED
    float l, r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"((float)x));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(l * -0.33333334f));
    return (double)r;
#else
    return (double)rcbrtf((float)x);
#endif
}

__device__ __forceinline__ double pow_two_thirds(double x) {
#if TF_FAST_P23
    // x^(2/3) = x * x^(-1/3); r = float estimate of x^(-1/3) (relative error e
    // ~ 1e-7), h = 1 - x r^3 ~ 3e, x^(-1/3) = r (1 - h)^(-1/3) = r (1 + h/3 +
    // 2h^2/9 + O(h^3)): a few ulp of float64, far inside the 1e-10 bar on u
    // (the reference's pow is itself ~1 ulp off the exact value).
    if (x > 1e-30 && x < 1e30) {
        const double r = rcbrt_seed(x);
        const double xr = x * r;
        const double h = __fma_rn(-(xr * r), r, 1.0);
        const double corr = __fma_rn(h, __fma_rn(h, 2.0 / 9.0, 1.0 / 3.0), 1.0);
        return xr * corr;
    }
#endif
    const double c = cbrt(x);
    return c * c;
}

// S_free = S_bed + (d - d_parent) / ds (channels_base.py:1542-1543) inside the step kernels, where
// it only feeds the velocity; get_values re-derives it with the exact division (k_derive)
// (exact = true for the dynamic wave: its momentum balance amplifies last-bit differences)
__device__ __forceinline__ double free_slope_step(double S_bed, double d, double d_parent, double ds, bool exact) {
#if TF_FAST_MANNING
    if (!exact) return S_bed + div_approx(d - d_parent, ds);
#endif
    return S_bed + ((d - d_parent) / ds);
}

// manning_formula :2561-2589 / law_of_the_wall :2593-2640
template <int LAW>
__device__ __forceinline__ double chan_velocity(double Rh, double S, double rough, double d) {
    if (LAW == 0) {
#if TF_FAST_MANNING
        return div_approx(pow_two_thirds(Rh) * sqrt(S), rough);    // (u only: within ~2 ulp of the quotient)
#else
        return (pow_two_thirds(Rh) * sqrt(S)) / rough;
#endif
    } else {
        double sm = (TF_AVAL / rough) * d;
        sm = np_maxv(sm, 1.1);
        const double law_const = sqrt(TF_G) / TF_KAPPA;     // :459
        return (law_const * sqrt(Rh * S)) * log(sm);
    }
}

// x ** (-1/3) for x > 0, same construction as pow_two_thirds
__device__ __forceinline__ double inv_cbrt(double x) {
#if TF_FAST_P23
    if (x > 1e-30 && x < 1e30) {
        const double r = rcbrt_seed(x);
        const double h = __fma_rn(-((x * r) * r), r, 1.0);
        return r * __fma_rn(h, __fma_rn(h, 2.0 / 9.0, 1.0 / 3.0), 1.0);
    }
#endif
    return 1.0 / cbrt(x);
}

// update_friction_factor :1654-1721
template <int LAW>
__device__ __forceinline__ double chan_friction(double d, double rough) {
    if (!(d > 0.0)) return 0.0;
    if (LAW == 0) {
#if TF_FAST_P23
        return TF_G * ((rough * rough) * inv_cbrt(d));      // g n^2 / d^(1/3), a few ulp
#else
        return TF_G * ((rough * rough) / cbrt(d));
#endif
    } else {
        double sm = (TF_AVAL / rough) * d;
        sm = np_maxv(sm, 1.1);
        const double r = TF_KAPPA / log(sm);
        return r * r;
    }
}

__device__ __forceinline__ bool bad_value(double v) {
    return (v < 0.0) || !isfinite(v);                        // :1992, :2044
}

// ---- CPT-wide row segment loads / stores (CPT = cells per thread) --------------
#ifndef TF_CPT
#define TF_CPT 2
#endif
#define CPT TF_CPT
static_assert(CPT == 1 || CPT == 2 || CPT == 4, "cells per thread must be 1, 2 or 4");

template <typename V, typename T, bool NC>
__device__ __forceinline__ V ldv(const T* p) {
    if (NC) return __ldg(reinterpret_cast<const V*>(p));
    return *reinterpret_cast<const V*>(p);
}

template <bool NC>
__device__ __forceinline__ void ld_f64(const double* p, double (&o)[CPT]) {
#if TF_CPT == 1
    o[0] = ldv<double, double, NC>(p);
#else
    const double2 a = ldv<double2, double, NC>(p);
    o[0] = a.x; o[1] = a.y;
#if TF_CPT == 4
    const double2 b = ldv<double2, double, NC>(p + 2);
    o[2] = b.x; o[3] = b.y;
#endif
#endif
}
template <bool NC>
__device__ __forceinline__ void ld_u8(const uint8_t* p, int (&o)[CPT]) {
#if TF_CPT == 1
    o[0] = ldv<unsigned char, uint8_t, NC>(p);
#elif TF_CPT == 2
    const uchar2 a = ldv<uchar2, uint8_t, NC>(p);
    o[0] = a.x; o[1] = a.y;
#else
    const uchar4 a = ldv<uchar4, uint8_t, NC>(p);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w;
#endif
}
// read-only (non-coherent) path: rasters no kernel of this step writes
__device__ __forceinline__ void ld4(const double* __restrict__ p, double (&o)[CPT]) { ld_f64<true>(p, o); }
__device__ __forceinline__ void ld4(const uint8_t* __restrict__ p, int (&o)[CPT]) { ld_u8<true>(p, o); }
__device__ __forceinline__ void ld4(const float* __restrict__ p, double (&o)[CPT]) {
#if TF_CPT == 1
    o[0] = __ldg(p);
#elif TF_CPT == 2
    const float2 a = __ldg(reinterpret_cast<const float2*>(p));
    o[0] = a.x; o[1] = a.y;
#else
    const float4 a = __ldg(reinterpret_cast<const float4*>(p));
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w;
#endif
}
// plain (coherent) loads for buffers this kernel, or an earlier kernel of the
// same step, writes
__device__ __forceinline__ void ld4c(const double* p, double (&o)[CPT]) { ld_f64<false>(p, o); }
__device__ __forceinline__ void ld4c(const uint8_t* p, int (&o)[CPT]) { ld_u8<false>(p, o); }

__device__ __forceinline__ void st4(double* p, const double (&v)[CPT]) {
#if TF_CPT == 1
    p[0] = v[0];
#else
    reinterpret_cast<double2*>(p)[0] = make_double2(v[0], v[1]);
#if TF_CPT == 4
    reinterpret_cast<double2*>(p)[1] = make_double2(v[2], v[3]);
#endif
#endif
}
__device__ __forceinline__ void st4(uint8_t* p, const int (&v)[CPT]) {
#if TF_CPT == 1
    p[0] = (uint8_t)v[0];
#elif TF_CPT == 2
    *reinterpret_cast<uchar2*>(p) = make_uchar2(v[0], v[1]);
#else
    *reinterpret_cast<uchar4*>(p) = make_uchar4(v[0], v[1], v[2], v[3]);
#endif
}
