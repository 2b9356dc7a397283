// tf_chan_march.cuh -- the routing step as a marching kernel (packed layout).
//
// One warp owns a band of 64 consecutive columns and walks down a segment of
// rows.  Everything it reads arrives through cp.async (global -> shared, no
// registers held) D rows ahead of the row being computed, so DRAM latency is
// covered by the warp's own arithmetic instead of by occupancy, addresses are
// a handful of per-lane pointers that advance by one row per iteration, and
// the eight-neighbour inflow is gathered from the three discharge rows the
// warp keeps in shared memory.
//
// Kinematic wave (WAVE 0): per row, volume -> depth -> velocity -> next
// discharge, as k_step_kinematic.
//
// Diffusive / dynamic wave (WAVE 1, 2): the free-surface slope of a cell needs
// the NEW depth of its parent (channels_base.py:1536-1543), which the tile
// kernels get from a second kernel after a grid-wide barrier.  Here the warp
// computes volume and depth of row y + 1 ("phase A"), keeps the last three
// rows of new depths in shared memory, and then finishes row y ("phase B":
// slope, velocity, next discharge).  Bands overlap by two columns on each side
// (lanes 0 and 31 only recompute their neighbours' cells) and segments by one
// row above and two below, so no warp ever waits for another.  What the
// recomputation reads must still hold its start-of-step value when a
// neighbour has already finished: the volume is double-buffered and the
// overbank byte keeps two generations (mstate_fired).
//
// A row is "phase A'd" by its owner and, redundantly, by the neighbouring
// segment / band; only the owner stores it.  A strip with a neighbouring strip
// (tf_set_strip) also phase-A's the neighbour's edge row into its own apron
// row, from the two apron rows of discharge it received: one exchange per
// step, at the end, for every wave.
#pragma once
#include "tf_chan_kernels2.cuh"
#include "tf_common.cuh"

#ifndef TF_MARCH_D
#define TF_MARCH_D 2            /* rows in flight ahead of the row being computed */
#endif
// One warp per block: everything that depends on the task only (rows, ring
// positions, the base addresses of the wide copies) is uniform across the
// block, so the compiler keeps it in uniform registers and the per-thread
// register file holds cell data only.
#ifndef TF_MARCH_UNROLL_K
#define TF_MARCH_UNROLL_K 2     /* 2: both cells of a lane interleaved; 1: one after the other */
#endif
static constexpr int kMarchUnrollK = TF_MARCH_UNROLL_K;
#ifndef TF_MARCH_MINB
#define TF_MARCH_MINB 12        /* blocks (= warps) per SM the fused waves are compiled for */
#endif
#ifndef TF_MARCH_MINB_KIN
#define TF_MARCH_MINB_KIN 16    /* ... and the kinematic wave */
#endif

// ---------------------------------------------------------------------------
// Per-warp shared-memory layout.  Streams are grouped in classes; all streams
// of a class hold the same rows, and a class slot keeps one row of each.
//   class 0: Q                      rows yA-2 (dynamic) / yA-1 .. yA+1 + D ahead
//   class 1: vol, w, code, inmask, mstate, aidx, row lengths, da, [R], [kfac]
//            (and S, n when phase B runs in the same iteration: WAVE 0, LAW 1)
//                                   row yA + D ahead
//   class 2: S, n (WAVE != 0)       row yA-1 (phase B) + D ahead
//   class 3: u_old (dynamic)        rows yA-2 .. yA + D ahead
//   d ring : new depths, rows yA-2 .. yA (written by the warp itself)
// ---------------------------------------------------------------------------
template <int WAVE, int LAW, bool RGRID>
struct MarchCfg {
    static constexpr bool HALO = (WAVE != 0);
    static constexpr bool KFAC = (WAVE == 0 && LAW == 0);
    static constexpr bool SN_C1 = (WAVE == 0 && LAW != 0);   // S, n needed with phase A's row
    static constexpr bool SN_C2 = HALO;
    static constexpr bool DYN = (WAVE == 2);
    static constexpr bool RGRID_ = RGRID;
    static constexpr int WAVE_ = WAVE, LAW_ = LAW;
    static constexpr int D = TF_MARCH_D;
    static constexpr int STRIDE = HALO ? 60 : 64;      // columns a band owns
    static constexpr int S0 = HALO ? -2 : 0;           // first window column of band 0
    static constexpr int FOFF = HALO ? 2 : 0;          // float windows start at s - FOFF
    static constexpr int FCH = HALO ? 17 : 16;         // 16-byte chunks of a float window
    static constexpr int BCH = HALO ? 5 : 4;           // 16-byte chunks of a byte window
    // ring depths (rows)
    static constexpr int R0 = D + (DYN ? 4 : 3);
    static constexpr int R1 = D + 1;
    static constexpr int R2 = D + 1;
    static constexpr int R3 = D + 3;
    // class 0 slot
    static constexpr int QB = 68 * 8;
    // class 1 slot: byte offsets of each stream inside the slot
    static constexpr int O_VOL = 0;
    static constexpr int O_W = O_VOL + 512;
    static constexpr int O_CODE = O_W + FCH * 16;
    static constexpr int O_MASK = O_CODE + BCH * 16;
    static constexpr int O_MST = O_MASK + BCH * 16;
    static constexpr int O_AIDX = O_MST + BCH * 16;
    static constexpr int O_LEN = O_AIDX + BCH * 16;            // 8 floats + 4 doubles (TF_LEN_ROW)
    static constexpr int O_DA = O_LEN + 64;                    // 1 double (+ 8 pad)
    static constexpr int O_RG = O_DA + 16;
    static constexpr int O_KF = O_RG + (RGRID ? 512 : 0);
    static constexpr int O_S1 = O_KF + (KFAC ? 512 : 0);
    static constexpr int O_N1 = O_S1 + (SN_C1 ? FCH * 16 : 0);
    static constexpr int C1B = O_N1 + (SN_C1 ? FCH * 16 : 0);
    // class 2 slot
    static constexpr int C2B = SN_C2 ? 2 * FCH * 16 : 0;
    // class 3 slot
    static constexpr int UB = DYN ? 68 * 8 : 0;
    // rings inside the warp's region
    static constexpr int P_Q = 0;
    static constexpr int P_DN = P_Q + R0 * QB;                 // 3 rows x 64 doubles (+ 16 B pad each side)
    static constexpr int DNB = HALO ? (512 + 32) : 0;
    static constexpr int P_C1 = P_DN + 3 * DNB;
    static constexpr int P_C2 = P_C1 + R1 * C1B;
    static constexpr int P_C3 = P_C2 + R2 * C2B;
    static constexpr int WARP_BYTES = P_C3 + R3 * UB;
    // cp.async chunks per row and instructions per lane
    static constexpr int N_CHUNKS = 34 /*Q*/ + 32 /*vol*/ + FCH /*w*/ + 4 * BCH + 4 /*len*/ + 1 /*da*/
                                    + (RGRID ? 32 : 0) + (KFAC ? 32 : 0) + ((SN_C1 || SN_C2) ? 2 * FCH : 0)
                                    + (DYN ? 34 : 0);
    // the short streams share KT per-lane copies; streams of exactly 32 chunks
    // are "wide" copies with a uniform base address
    static constexpr int N_TAIL = 2 + (DYN ? 2 : 0) + FCH + 4 * BCH + 4 + 1 + ((SN_C1 || SN_C2) ? 2 * FCH : 0);
    static constexpr int KT = (N_TAIL + 31) / 32;
    static constexpr size_t SMEM = WARP_BYTES;
};

__device__ __forceinline__ unsigned smem_u32(const void* p) {
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void cp_async16(unsigned dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(unsigned dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// one cp.async a lane issues per row: where it reads, where it lands
struct MarchCopy {
    const char* g;      // global address of this lane's chunk for the next row to issue
    int stride;         // bytes from one row to the next
    int dst;            // shared-memory byte offset inside the class slot (+ ring base)
    int cls;            // 0..3, -1: nothing to do; bit 4: an 8-byte copy
};


// Which chunk of which short stream does tail copy number c (0 .. 32 KT - 1)
// of a band starting at column s move?  Addresses are for row 0; the caller
// offsets them.  (The streams of exactly 32 chunks -- Q's first 512 bytes, vol,
// R, kfac, u_old -- are issued with uniform base addresses in k_march.)
template <typename C>
__device__ __forceinline__ MarchCopy march_assign(const ChanArgs& a, int s, int c) {
    MarchCopy r;
    r.g = nullptr; r.stride = 0; r.dst = 0; r.cls = -1;
    const int p = a.pitch;
    const int sb = s & ~15;                        // byte windows start 16-aligned
    // ---- the short streams share the remaining copies ----
    if (c < 2) { r.g = (const char*)(a.Q + (s - 2)) + (32 + c) * 16; r.stride = p * 8; r.dst = C::P_Q + (32 + c) * 16; r.cls = 0; return r; }
    c -= 2;
    if (C::DYN) {
        if (c < 2) { r.g = (const char*)(a.u_old + (s - 2)) + (32 + c) * 16; r.stride = p * 8; r.dst = C::P_C3 + (32 + c) * 16; r.cls = 3; return r; }
        c -= 2;
    }
    if (c < C::FCH) { r.g = (const char*)(a.w32 + (s - C::FOFF)) + c * 16; r.stride = p * 4; r.dst = C::P_C1 + C::O_W + c * 16; r.cls = 1; return r; }
    c -= C::FCH;
    if (c < C::BCH) { r.g = (const char*)(a.code + sb) + c * 16; r.stride = p; r.dst = C::P_C1 + C::O_CODE + c * 16; r.cls = 1; return r; }
    c -= C::BCH;
    if (c < C::BCH) { r.g = (const char*)(a.inmask + sb) + c * 16; r.stride = p; r.dst = C::P_C1 + C::O_MASK + c * 16; r.cls = 1; return r; }
    c -= C::BCH;
    if (c < C::BCH) { r.g = (const char*)(a.mstate + sb) + c * 16; r.stride = p; r.dst = C::P_C1 + C::O_MST + c * 16; r.cls = 1; return r; }
    c -= C::BCH;
    if (c < C::BCH) { r.g = (const char*)(a.aidx + sb) + c * 16; r.stride = p; r.dst = C::P_C1 + C::O_AIDX + c * 16; r.cls = 1; return r; }
    c -= C::BCH;
    if (c < 4) { r.g = (const char*)a.len_rows + c * 16; r.stride = TF_LEN_ROW * 4; r.dst = C::P_C1 + C::O_LEN + c * 16; r.cls = 1; return r; }
    c -= 4;
    if (c < 1) { r.g = (const char*)a.da_rows; r.stride = 8; r.dst = C::P_C1 + C::O_DA; r.cls = 1 | 16; return r; }
    c -= 1;
    if (C::SN_C1 || C::SN_C2) {
        const int cls = C::SN_C1 ? 1 : 2;
        const int base = C::SN_C1 ? (C::P_C1 + C::O_S1) : C::P_C2;
        if (c < C::FCH) { r.g = (const char*)(a.S32 + (s - C::FOFF)) + c * 16; r.stride = p * 4; r.dst = base + c * 16; r.cls = cls; return r; }
        c -= C::FCH;
        if (c < C::FCH) { r.g = (const char*)(a.n32 + (s - C::FOFF)) + c * 16; r.stride = p * 4; r.dst = base + C::FCH * 16 + c * 16; r.cls = cls; return r; }
        c -= C::FCH;
    }
    return r;
}

// ring slot of a row, used once per task to start the running offsets
template <int R>
__device__ __forceinline__ int march_slot(int row) { return (row + 2 + 16 * R) % R; }

// a ring position (byte offset from the ring's start) one row further on
template <int R, int BYTES>
__device__ __forceinline__ int ring_next(int off) { return (off + BYTES == R * BYTES) ? 0 : off + BYTES; }

// what phase B needs from phase A of the same row, one iteration later
struct MarchCarry {
    double w[2], t[2], c[2], ds[2], d[2], v[2];
    int code[2], mask[2];
};

// direction j = 0..7 (code 1 << j) -> (row, column) step towards the parent
__device__ __forceinline__ int dir_dr(int j) { return (j == 0 || j >= 6) ? -1 : ((j >= 2 && j <= 4) ? 1 : 0); }
__device__ __forceinline__ int dir_dc(int j) { return (j <= 2) ? 1 : ((j >= 4 && j <= 6) ? -1 : 0); }

// acc += dt * q when bit `bit` of mask is set: one predicated multiply and add,
// no branch (channels_base.py:1346-1361, one term of the k = 1..8 scatter)
template <int BIT>
__device__ __forceinline__ void add_inflow(double& acc, int mask, double dt, double q) {
    asm("{\n\t.reg .pred p;\n\t.reg .b32 m;\n\t.reg .f64 t;\n\t"
        "and.b32 m, %1, %2;\n\tsetp.ne.b32 p, m, 0;\n\t"
        "@p mul.rn.f64 t, %3, %4;\n\t"
        "@p add.rn.f64 %0, %0, t;\n\t}"
        : "+d"(acc) : "r"(mask), "n"(BIT), "d"(dt), "d"(q));
}

template <int WAVE, int LAW, bool RGRID>
__global__ void __launch_bounds__(32, (WAVE == 0) ? TF_MARCH_MINB_KIN : TF_MARCH_MINB)
k_march(const __grid_constant__ ChanArgs a) {
    using C = MarchCfg<WAVE, LAW, RGRID>;
    constexpr int D = C::D;
    extern __shared__ __align__(16) unsigned char march_smem[];
    const int lane = threadIdx.x;                           // one warp per block
    const int band = (int)(blockIdx.x % (unsigned)a.n_bands);
    const int seg = (int)(blockIdx.x / (unsigned)a.n_bands);
    const int y0 = a.y_begin + seg * a.seg_rows;
    const int y1 = min(y0 + a.seg_rows, a.y_end);
    const int s = band * C::STRIDE + C::S0;                 // first column of the 64-wide window
    unsigned char* wsm = march_smem;
    const unsigned wsm_u = smem_u32(wsm);
    const int ny = a.ny, pitch = a.pitch;
    const double* __restrict__ lut = a.lut;

    // rows phase A visits
    const int yA_first = C::HALO ? y0 - 1 : y0;
    const int yA_last = C::HALO ? y1 : y1 - 1;
    const int i_start = yA_first - (D + 2);

    // ---- copies ---------------------------------------------------------------
    // wide streams: 32 chunks of 16 bytes per row, lane l moves chunk l; the
    // addresses below are uniform (lane 0's chunk of the next row to issue)
    const int64_t row_c1 = (int64_t)(i_start + D) * pitch + s;          // class 1 and 3 rows
    const char* gQ = reinterpret_cast<const char*>(a.Q + row_c1 + pitch - 2);       // class 0: one row further
    const char* gV = reinterpret_cast<const char*>(a.vol + row_c1);
    const char* gR = RGRID ? reinterpret_cast<const char*>(a.Rgrid + row_c1) : nullptr;
    const char* gK = C::KFAC ? reinterpret_cast<const char*>(a.kfac + row_c1) : nullptr;
    const char* gU = C::DYN ? reinterpret_cast<const char*>(a.u_old + row_c1 - 2) : nullptr;
    const int64_t stride8 = (int64_t)pitch * 8;
    const int lane16 = lane * 16;
    // short streams: per-lane copies
    MarchCopy cp[C::KT];
#pragma unroll
    for (int k = 0; k < C::KT; ++k) {
        cp[k] = march_assign<C>(a, s, k * 32 + lane);
        const int cls = cp[k].cls & 15;
        const int roff = (cls == 0) ? D + 1 : ((cls == 2) ? D - 1 : D);
        if (cp[k].cls >= 0) cp[k].g += (long long)(i_start + roff) * cp[k].stride;
    }
    // ring positions (byte offsets) the next issue writes to, per class
    int iq = march_slot<C::R0>(i_start + D + 1) * C::QB;
    int i1 = march_slot<C::R1>(i_start + D) * C::C1B;
    int i2 = march_slot<C::R2>(i_start + D - 1) * C::C2B;
    int i3 = march_slot<C::R3>(i_start + D) * C::UB;
    auto issue = [&](int i) {
        const int r0 = i + D + 1, r1 = i + D, r2 = i + D - 1;
        // Only rows that exist and that this task will read: a copy of a row
        // nobody needs would share its ring slot with a later, needed row while
        // both are still in flight, and the two may land in either order.
        const bool ok0 = r0 >= -TF_APRON && r0 < ny + TF_APRON;                      // first: yA_first - 1
        const bool ok1 = r1 >= yA_first && r1 < ny + TF_APRON;
        const bool ok2 = r2 >= y0 && r2 < ny + TF_APRON;
        const bool ok3 = r1 >= y0 - 1 && r1 >= -TF_APRON && r1 < ny + TF_APRON;
        if (ok0) cp_async16(wsm_u + (unsigned)(C::P_Q + iq + lane16), gQ + lane16);
        if (ok1) {
            cp_async16(wsm_u + (unsigned)(C::P_C1 + C::O_VOL + i1 + lane16), gV + lane16);
            if (RGRID) cp_async16(wsm_u + (unsigned)(C::P_C1 + C::O_RG + i1 + lane16), gR + lane16);
            if (C::KFAC) cp_async16(wsm_u + (unsigned)(C::P_C1 + C::O_KF + i1 + lane16), gK + lane16);
        }
        if (C::DYN && ok3) cp_async16(wsm_u + (unsigned)(C::P_C3 + i3 + lane16), gU + lane16);
        gQ += stride8; gV += stride8;
        if (RGRID) gR += stride8;
        if (C::KFAC) gK += stride8;
        if (C::DYN) gU += stride8;
#pragma unroll
        for (int k = 0; k < C::KT; ++k) {
            const int cls = cp[k].cls & 15;
            const bool ok = (cp[k].cls >= 0) && ((cls == 0) ? ok0 : ((cls == 1) ? ok1 : ((cls == 2) ? ok2 : ok3)));
            const int base = (cls == 0) ? iq : ((cls == 1) ? i1 : ((cls == 2) ? i2 : i3));
            if (ok) {
                const unsigned dst = wsm_u + (unsigned)(cp[k].dst + base);
                if (cp[k].cls & 16) cp_async8(dst, cp[k].g);
                else cp_async16(dst, cp[k].g);
            }
            cp[k].g += cp[k].stride;
        }
        cp_async_commit();
        iq = ring_next<C::R0, C::QB>(iq);
        i1 = ring_next<C::R1, C::C1B>(i1);
        if (C::SN_C2) i2 = ring_next<C::R2, C::SN_C2 ? C::C2B : 16>(i2);
        if (C::DYN) i3 = ring_next<C::R3, C::DYN ? C::UB : 16>(i3);
    };
    for (int i = i_start; i < yA_first; ++i) issue(i);

    // ring positions the computation reads: Q rows yA-2 .. yA+1, the class-1
    // row yA, S / n of row yA-1, u_old and new depths of rows yA-2 .. yA
    int oqmm = march_slot<C::R0>(yA_first - 2) * C::QB, oqm = march_slot<C::R0>(yA_first - 1) * C::QB;
    int oq0 = march_slot<C::R0>(yA_first) * C::QB, oqp = march_slot<C::R0>(yA_first + 1) * C::QB;
    int o1 = march_slot<C::R1>(yA_first) * C::C1B;
    int o2 = march_slot<C::R2>(yA_first - 1) * C::C2B;
    int oum = march_slot<C::R3>(yA_first - 2) * C::UB, ou0 = march_slot<C::R3>(yA_first - 1) * C::UB;
    int oup = march_slot<C::R3>(yA_first) * C::UB;
    int odm = 0, od0 = C::DNB, odp = 2 * C::DNB;

    const int g_in = (int)(a.steps_done & 1), g_out = g_in ^ 1;
    const bool own_lane = C::HALO ? (lane >= 1 && lane <= 30) : true;
    const int x = s + 2 * lane;                             // this lane's first column
    const bool in_row = own_lane && x < pitch;
    const int boff = s - (s & ~15);                         // byte windows: first column's offset
    const double dt = a.dt;
    const bool mut_w = a.mut_width != 0, mut_a = a.mut_angle != 0;
    MarchCarry cur;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        cur.w[k] = cur.t[k] = cur.c[k] = cur.ds[k] = cur.d[k] = cur.v[k] = 0.0;
        cur.code[k] = cur.mask[k] = 0;
    }

    for (int yA = yA_first; yA <= yA_last; ++yA) {
        issue(yA);
        cp_async_wait<D>();
        __syncwarp();
        MarchCarry nxt;
        // =====================================================================
        // phase A of row yA: update_flow_volume :1308-1375, update_flow_depth
        // :1379-1532
        // =====================================================================
        const bool rowA = !C::HALO || (yA >= 0 && yA < ny) || (yA == -1 && a.has_north) || (yA == ny && a.has_south);
        const bool ownA = (yA >= y0 && yA < y1);
        const bool storeA = ownA || (C::HALO && (yA == -1 || yA == ny) && rowA);
        double* dn_row = reinterpret_cast<double*>(wsm + C::P_DN + odp + 16);
        if (rowA) {
            const unsigned char* c1 = wsm + C::P_C1 + o1;
            const double* Qm = reinterpret_cast<const double*>(wsm + C::P_Q + oqm);
            const double* Q0 = reinterpret_cast<const double*>(wsm + C::P_Q + oq0);
            const double* Qp = reinterpret_cast<const double*>(wsm + C::P_Q + oqp);
            const int qi = 2 + 2 * lane;
            double qm[4], q0[4], qp[4];                     // window columns qi-1 .. qi+2 of the three rows
            {
                qm[0] = Qm[qi - 1]; const double2 t2 = *reinterpret_cast<const double2*>(Qm + qi);
                qm[1] = t2.x; qm[2] = t2.y; qm[3] = Qm[qi + 2];
            }
            {
                q0[0] = Q0[qi - 1]; const double2 t2 = *reinterpret_cast<const double2*>(Q0 + qi);
                q0[1] = t2.x; q0[2] = t2.y; q0[3] = Q0[qi + 2];
            }
            {
                qp[0] = Qp[qi - 1]; const double2 t2 = *reinterpret_cast<const double2*>(Qp + qi);
                qp[1] = t2.x; qp[2] = t2.y; qp[3] = Qp[qi + 2];
            }
            const int bi = boff + 2 * lane;
            double accv[2];
            int maskv[2];
            {
                // ---- stage 1: mass balance (only what it needs is loaded yet) ----
                const double2 v2 = *reinterpret_cast<const double2*>(c1 + C::O_VOL + lane * 16);
                const uchar2 mask2 = *reinterpret_cast<const uchar2*>(c1 + C::O_MASK + bi);
                const double da = *reinterpret_cast<const double*>(c1 + C::O_DA);
                double R[2];
                if (RGRID) {
                    const double2 r2 = *reinterpret_cast<const double2*>(c1 + C::O_RG + lane * 16);
                    R[0] = r2.x; R[1] = r2.y;
                } else { R[0] = R[1] = a.Rs; }
                const double vin[2] = {v2.x, v2.y};
                maskv[0] = mask2.x; maskv[1] = mask2.y;
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    double acc = vin[k] + (R[k] * da) * dt;                // :1334
                    const int mk = maskv[k];
                    // the k = 1..8 adds of :1346-1361, in order, each on its mask bit
                    add_inflow<1>(acc, mk, dt, qp[k]);                     // (+1, -1)
                    add_inflow<2>(acc, mk, dt, q0[k]);                     // ( 0, -1)
                    add_inflow<4>(acc, mk, dt, qm[k]);                     // (-1, -1)
                    add_inflow<8>(acc, mk, dt, qm[k + 1]);                 // (-1,  0)
                    add_inflow<16>(acc, mk, dt, qm[k + 2]);                // (-1, +1)
                    add_inflow<32>(acc, mk, dt, q0[k + 2]);                // ( 0, +1)
                    add_inflow<64>(acc, mk, dt, qp[k + 2]);                // (+1, +1)
                    add_inflow<128>(acc, mk, dt, qp[k + 1]);               // (+1,  0)
                    acc -= q0[k + 1] * dt;                                 // :1366
                    accv[k] = np_max0(acc);                                // :1375
                }
            }
            // keep the compiler from hoisting the next stage's loads above the
            // gather: the three discharge rows are dead by now, and a warp that
            // holds fewer values at once lets more warps share the SM
            asm volatile("" ::: "memory");
            // ---- stage 2: trapezoid depth and the overbank rule ----
            const float2 wf = *reinterpret_cast<const float2*>(c1 + C::O_W + (C::FOFF + 2 * lane) * 4);
            const uchar2 code2 = *reinterpret_cast<const uchar2*>(c1 + C::O_CODE + bi);
            const uchar2 mst2 = *reinterpret_cast<const uchar2*>(c1 + C::O_MST + bi);
            const uchar2 aidx2 = *reinterpret_cast<const uchar2*>(c1 + C::O_AIDX + bi);
            const float4 rs = *reinterpret_cast<const float4*>(c1 + C::O_LEN);
            const float4 rw = *reinterpret_cast<const float4*>(c1 + C::O_LEN + 16);
            const double2 rcp01 = *reinterpret_cast<const double2*>(c1 + C::O_LEN + 32);   // 1/ds: diag, E/W
            const double rcp2 = *reinterpret_cast<const double*>(c1 + C::O_LEN + 48);      //       N/S
            const double wv[2] = {(double)wf.x, (double)wf.y};
            const int codev[2] = {code2.x, code2.y};
            const int mstv[2] = {mst2.x, mst2.y}, aiv[2] = {aidx2.x, aidx2.y};
            int newb[2];
            bool changed = false;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int cd = codev[k];
                const int gin = (mstv[k] >> (2 * g_in)) & 3;
                const int m = (gin | (gin >> 1)) & 1;
                double w = wv[k], t = __ldg(lut + aiv[k]), c = __ldg(lut + 256 + aiv[k]);
                const double rcp_den = __ldg(lut + 512 + aiv[k]);
                const bool diag = (cd & 0x55) != 0;
                const double ds = (double)(diag ? rs.x : ((cd & 0x88) ? rs.z : rs.y));
                const double rcp_ds = diag ? rcp01.x : ((cd & 0x88) ? rcp2 : rcp01.y);
                const double dw = (double)(diag ? rw.x : ((cd & 0x22) ? rw.y : rw.z));
                if (m) {                                               // overbank rule has fired (:1449-1454)
                    if (mut_w) w = dw;
                    if (mut_a) { t = 0.0; c = 1.0; }
                }
                const double acc = accv[k];
                double dd = chan_depth_rcp(acc, w, t, ds, rcp_ds, rcp_den);
                if (cd == 0) dd = 0.0;                                 // :1511
                dd = np_max0(dd);                                      // :1526
                // next step's overbank test with this step's depth (:1449-1450)
                const double top = w + ((2.0 * dd) * t);
                const int nb = m | ((top > dw) ? 2 : 0);
                changed = changed || (nb != ((mstv[k] >> (2 * g_out)) & 3));
                newb[k] = (mstv[k] & ~(3 << (2 * g_out))) | (nb << (2 * g_out));
                nxt.w[k] = w; nxt.t[k] = t; nxt.c[k] = c; nxt.ds[k] = ds; nxt.d[k] = dd; nxt.v[k] = acc;
                nxt.code[k] = cd; nxt.mask[k] = maskv[k];
            }
            if (C::HALO) *reinterpret_cast<double2*>(dn_row + 2 * lane) = make_double2(nxt.d[0], nxt.d[1]);
            if (storeA && in_row) {
                const int64_t i = (int64_t)yA * pitch + x;
                *reinterpret_cast<double2*>(a.vol_out + i) = make_double2(nxt.v[0], nxt.v[1]);
                *reinterpret_cast<double2*>(a.d + i) = make_double2(nxt.d[0], nxt.d[1]);
                if (changed) *reinterpret_cast<uchar2*>(a.mstate + i) = make_uchar2(newb[0], newb[1]);
                // d = np_max0(.) is never negative, so "bad" is NaN or +inf
                if (ownA && !(nxt.d[0] <= 1.7976931348623157e308 && nxt.d[1] <= 1.7976931348623157e308)) {
#pragma unroll
                    for (int k = 1; k >= 0; --k)
                        if (x + k < a.nx && bad_value(nxt.d[k])) flag_bad(a, 0, x + k, yA);
                }
            }
            if (!C::HALO) {
                // ---- kinematic wave: velocity and next discharge of the same row ----
                double S[2] = {0.0, 0.0}, n[2] = {1.0, 1.0};
                if (C::KFAC) {
                    const double2 kf = *reinterpret_cast<const double2*>(c1 + C::O_KF + lane * 16);
                    S[0] = kf.x; S[1] = kf.y;
                } else {
                    const float2 sf = *reinterpret_cast<const float2*>(c1 + C::O_S1 + (C::FOFF + 2 * lane) * 4);
                    const float2 nf = *reinterpret_cast<const float2*>(c1 + C::O_N1 + (C::FOFF + 2 * lane) * 4);
                    S[0] = sf.x; S[1] = sf.y; n[0] = nf.x; n[1] = nf.y;
                }
                double u[2], Qn[2];
#pragma unroll kMarchUnrollK
                for (int k = 0; k < 2; ++k) {
                    const bool noflow = (nxt.code[k] == 0);
                    const XSection xs = chan_xsection_step(nxt.d[k], nxt.w[k], nxt.t[k], nxt.c[k], noflow);
                    double uu = C::KFAC ? pow_two_thirds(xs.Rh) * S[k]
                                        : chan_velocity<LAW>(xs.Rh, S[k], n[k], nxt.d[k]);
                    if (noflow) uu = 0.0;                              // :1783-1791
                    u[k] = uu;
                    Qn[k] = uu * xs.A_wet;                             // :1188 of the next step
                }
                if (in_row) {
                    const int64_t i = (int64_t)yA * pitch + x;
                    *reinterpret_cast<double2*>(a.u + i) = make_double2(u[0], u[1]);
                    *reinterpret_cast<double2*>(a.Qn + i) = make_double2(Qn[0], Qn[1]);
                    if (!(u[0] >= 0.0 && u[0] <= 1.7976931348623157e308 && u[1] >= 0.0 && u[1] <= 1.7976931348623157e308)) {
#pragma unroll
                        for (int k = 1; k >= 0; --k)
                            if (x + k < a.nx && bad_value(u[k])) flag_bad(a, 1, x + k, yA);
                    }
                }
            }
        } else if (C::HALO) {
            // no such row (beyond a border of the global grid): nothing flows there
            *reinterpret_cast<double2*>(dn_row + 2 * lane) = make_double2(0.0, 0.0);
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                nxt.w[k] = 1.0; nxt.t[k] = 0.0; nxt.c[k] = 1.0; nxt.ds[k] = 1.0; nxt.d[k] = 0.0; nxt.v[k] = 0.0;
                nxt.code[k] = 0; nxt.mask[k] = 0;
            }
        }
        // =====================================================================
        // phase B of row yA - 1, whose parents' new depths are now all known
        // =====================================================================
        if (C::HALO) {
            __syncwarp();                                   // the depths of row yA are in shared memory
            const int yB = yA - 1;
            if (yB >= y0) {
                const unsigned char* c2 = wsm + C::P_C2 + o2;
                const float2 sf = *reinterpret_cast<const float2*>(c2 + (C::FOFF + 2 * lane) * 4);
                const float2 nf = *reinterpret_cast<const float2*>(c2 + C::FCH * 16 + (C::FOFF + 2 * lane) * 4);
                const double Sv[2] = {(double)sf.x, (double)sf.y}, nv[2] = {(double)nf.x, (double)nf.y};
                // rows yB-1, yB, yB+1 of the new depths, of Q and of u_old
                const double* dnr[3] = {reinterpret_cast<const double*>(wsm + C::P_DN + odm + 16),
                                        reinterpret_cast<const double*>(wsm + C::P_DN + od0 + 16),
                                        reinterpret_cast<const double*>(wsm + C::P_DN + odp + 16)};
                const double* qr[3] = {reinterpret_cast<const double*>(wsm + C::P_Q + oqmm),
                                       reinterpret_cast<const double*>(wsm + C::P_Q + oqm),
                                       reinterpret_cast<const double*>(wsm + C::P_Q + oq0)};
                const double* ur[3] = {reinterpret_cast<const double*>(wsm + C::P_C3 + oum),
                                       reinterpret_cast<const double*>(wsm + C::P_C3 + ou0),
                                       reinterpret_cast<const double*>(wsm + C::P_C3 + oup)};
                double u[2], Qn[2];
#pragma unroll kMarchUnrollK
                for (int k = 0; k < 2; ++k) {
                    const int idx = 2 * lane + k;
                    const int cd = cur.code[k];
                    const bool noflow = (cd == 0);
                    const double d = cur.d[k], w = cur.w[k], t = cur.t[k], c = cur.c[k], ds = cur.ds[k];
                    // free-surface slope, :1536-1543; a no-flow cell's parent is cell (0,0), d = 0
                    double dpar = 0.0;
                    if (cd) {
                        const int j = __ffs(cd) - 1;
                        const int dr = dir_dr(j), dc = dir_dc(j);
                        const double* row = (dr < 0) ? dnr[0] : ((dr > 0) ? dnr[2] : dnr[1]);
                        dpar = row[idx + dc];
                    }
                    double s_free = free_slope_step(Sv[k], d, dpar, ds, WAVE == 2);
                    double uu;
                    XSection xs;
                    if (WAVE == 1) {
                        xs = chan_xsection_step(d, w, t, c, noflow);
                        s_free = np_max0(s_free);                      // channels_diffusive_wave.py:118
                        uu = chan_velocity<LAW>(xs.Rh, s_free, nv[k], d);
                    } else {
                        // momentum balance, channels_dynamic_wave.py:126-341
                        xs = chan_xsection(d, w, t, c, noflow);
                        const double v = cur.v[k];
                        const double Qo = qr[1][2 + idx];
                        const double uo = ur[1][2 + idx];
                        const double u0 = (d > 0.0) ? uo : 0.0;        // :126-127
                        const double vinv = (v > 0.0) ? (1.0 / v) : 0.0;   // :140-145
                        const double Ap = ds * xs.P_wet;               // :166
                        const double Qtrm = u0 * Qo;                   // :171
                        const double grav = (TF_G * s_free) * v;       // :178
                        const double f = chan_friction<LAW>(d, nv[k]);
                        const double fric = ((f * u0) * u0) * Ap;      // :190
                        const double dM = ((grav - fric) - Qtrm) * dt; // :214
                        double M = (u0 * v) + dM;                      // :222
                        int mk = cur.mask[k];
                        while (mk) {                                   // :244-259, children in k = 1..8 order
                            const int j = __ffs(mk) - 1;
                            mk &= mk - 1;
                            const int dr = -dir_dr(j), dc = -dir_dc(j);   // the child sits opposite
                            const int rr = dr + 1;
                            const double* drow = (rr == 0) ? dnr[0] : ((rr == 2) ? dnr[2] : dnr[1]);
                            const double* qrow = (rr == 0) ? qr[0] : ((rr == 2) ? qr[2] : qr[1]);
                            const double* urow = (rr == 0) ? ur[0] : ((rr == 2) ? ur[2] : ur[1]);
                            const double uc = (drow[idx + dc] > 0.0) ? urow[2 + idx + dc] : 0.0;
                            M += (uc * qrow[2 + idx + dc]) * dt;
                        }
                        M = np_max0(M);                                // :312
                        uu = np_max0(M * vinv);                        // :318-331
                    }
                    if (noflow) uu = 0.0;                              // channels_base.py:1783-1791
                    u[k] = uu;
                    Qn[k] = uu * xs.A_wet;
                }
                if (in_row) {
                    const int64_t i = (int64_t)yB * pitch + x;
                    *reinterpret_cast<double2*>(a.u + i) = make_double2(u[0], u[1]);
                    *reinterpret_cast<double2*>(a.Qn + i) = make_double2(Qn[0], Qn[1]);
                    if (!(u[0] >= 0.0 && u[0] <= 1.7976931348623157e308 && u[1] >= 0.0 && u[1] <= 1.7976931348623157e308)) {
#pragma unroll
                        for (int k = 1; k >= 0; --k)
                            if (x + k < a.nx && bad_value(u[k])) flag_bad(a, 1, x + k, yB);
                    }
                }
            }
            cur = nxt;
            // the next iteration's rows
            { const int t3 = odm; odm = od0; od0 = odp; odp = t3; }
            o2 = ring_next<C::R2, C::SN_C2 ? C::C2B : 16>(o2);
            if (C::DYN) { oum = ou0; ou0 = oup; oup = ring_next<C::R3, C::DYN ? C::UB : 16>(oup); }
        }
        oqmm = oqm; oqm = oq0; oq0 = oqp; oqp = ring_next<C::R0, C::QB>(oqp);
        o1 = ring_next<C::R1, C::C1B>(o1);
        __syncwarp();                                       // every read of this iteration precedes the next issue
    }
    cp_async_wait<0>();
}
