// tf_d8.cu -- D8 flow-grid builder on the GPU.
//
// What the reference computes with ~40 whole-grid NumPy passes per call
// (topoflow/components/d8_global.py) is done here with a handful of kernels:
//   k_start_codes   slopes to the 8 neighbours in float32, steepest-descent
//                   code with ties OR-ed, flats / pits / nodata / borders, and
//                   the tie-break table         (d8_global.py:179-367)
//   k_link_flats    one Jacobi sweep of link_flats()   (d8_global.py:371-567)
//   k_finalize      uint8 codes, parent IDs, ds, dw    (:141-144, :574-644,
//                                                       :863-1004)
//   k_inmask        per-cell byte of "which neighbours drain into me", the
//                   transpose of the code grid; it replaces the reference's
//                   w1..w8 / p1..p8 index lists        (:719-835)
//   k_area_*        contributing area by dependency counting (:1018-1261)
// All integer/float32 results are bit-identical to the reference's.
#include <stdlib.h>

#include "tf_common.cuh"

// neighbour k = 0..7 : UR, R, LR, B, LL, L, UL, T  (d8_global.py:381-385)
__constant__ int c_dr[8] = {-1, 0, 1, 1, 1, 0, -1, -1};
__constant__ int c_dc[8] = {1, 1, 1, 0, -1, -1, -1, 0};
__constant__ uint8_t c_resolve[256];

#define PIT_CODE (-300)

// ---------------------------------------------------------------------------
// host: tie-break table (d8_base.py:744-816), built from its symmetry rules
// ---------------------------------------------------------------------------
static int rot45(int v, int times) {
    for (int i = 0; i < times; ++i) {
        v *= 2;
        if (v > 255) v += 1;
        v &= 255;
    }
    return v;
}

static void make_resolve(uint8_t* t) {
    memset(t, 0, 256);
    struct Grp { int value; int n; int m[8]; };
    static const Grp groups[] = {
        {1, 2, {17, 85}},  {2, 5, {34, 102, 119, 238, 170}},
        {8, 5, {136, 153, 187, 221, 255}}, {4, 1, {68}}, {32, 1, {51}},
        {128, 1, {204}},   {1, 3, {1, 5, 69}}, {4, 3, {4, 20, 21}},
        {16, 3, {16, 80, 84}}, {64, 3, {64, 65, 81}}};
    for (const Grp& g : groups)
        for (int i = 0; i < g.n; ++i) t[g.m[i]] = (uint8_t)g.value;
    static const int fam[] = {
        2, 3, 6, 7, 10, 11, 14, 15, 18, 19, 22, 23, 27, 31, 35, 38, 39, 43, 47,
        54, 55, 66, 67, 70, 71, 74, 75, 78, 79, 82, 83, 86, 87, 91, 95, 103,
        107, 111, 134, 135, 138, 139, 142, 143, 150, 151, 155, 159, 166, 167,
        171, 175, 183, 191, 206, 207, 223};
    const int nf = sizeof(fam) / sizeof(fam[0]);
    // same assignment order as the reference: later families overwrite
    for (int i = 0; i < nf; ++i) t[fam[i]] = 2;
    for (int i = 0; i < nf; ++i) t[rot45(fam[i], 2)] = 8;
    for (int i = 0; i < nf; ++i) t[rot45(fam[i], 4)] = 32;
    for (int i = 0; i < nf; ++i) t[rot45(fam[i], 6)] = 128;
}

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
struct D8Args {
    int nx, ny, pitch;
    int row_offset, global_ny;
    int has_north, has_south;
    int lr_periodic, tb_periodic;
};

__device__ __forceinline__ int wrap_x(int x, int nx) {
    return x < 0 ? x + nx : (x >= nx ? x - nx : x);
}

// np.maximum: NaN if either operand is NaN
__device__ __forceinline__ float np_max(float a, float b) {
    return (a != a) ? a : ((b != b) ? b : (a > b ? a : b));
}

// ---------------------------------------------------------------------------
// start_new_d8_codes + break_flow_grid_ties
//   dem aprons hold the wrapped / neighbour rows (np.roll semantics).
//   counters[0..2] += flats, pits, nodata
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_start_codes(D8Args a, const float* __restrict__ dem, int16_t* __restrict__ raw,
              float dx0, float dy0, float dd0, float nodata, int link,
              unsigned long long* counters) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    unsigned n_flat = 0, n_pit = 0, n_bad = 0;
    if (x < a.nx && y < a.ny) {
        const int64_t row = (int64_t)y * a.pitch;
        const float z = dem[row + x];
        const float dist[8] = {dd0, dx0, dd0, dy0, dd0, dx0, dd0, dy0};
        float s[8];
        float smax;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            int xx = wrap_x(x + c_dc[k], a.nx);
            float zn = dem[row + (int64_t)c_dr[k] * a.pitch + xx];
            s[k] = (z - zn) / dist[k];                        // d8_global.py:225-232
        }
        smax = np_max(s[0], s[1]);
#pragma unroll
        for (int k = 2; k < 8; ++k) smax = np_max(smax, s[k]);   // :250-256
        int code = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) code += (s[k] == smax) ? (1 << k) : 0;   // :266-274
        if (!link) {
            if (smax <= 0.0f) code = 0;                        // :281-285
        } else {
            if (smax == 0.0f) { code = -code; n_flat = 1; }    // :290-293
            if (smax < 0.0f) { code = PIT_CODE; n_pit = 1; }   // :299-302
        }
        if (z <= nodata || !isfinite(z)) { code = 0; n_bad = 1; }   // :312-315
        const int gy = y + a.row_offset;
        if (!a.lr_periodic && (x == 0 || x == a.nx - 1)) code = 0;   // :320-322
        if (!a.tb_periodic && (gy == 0 || gy == a.global_ny - 1)) code = 0;   // :327-329
        if (code > 0) code = c_resolve[code];                  // :365-367
        raw[row + x] = (int16_t)code;
    }
    // block-level counting: ballot per warp, one atomic per warp
    unsigned f = __ballot_sync(0xffffffffu, n_flat);
    unsigned p = __ballot_sync(0xffffffffu, n_pit);
    unsigned b = __ballot_sync(0xffffffffu, n_bad);
    if (((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0) {
        if (f) atomicAdd(&counters[0], (unsigned long long)__popc(f));
        if (p) atomicAdd(&counters[1], (unsigned long long)__popc(p));
        if (b) atomicAdd(&counters[2], (unsigned long long)__popc(b));
    }
}

// ---------------------------------------------------------------------------
// link_flats, one Jacobi sweep (reads `in`, writes `out`)
//   counters[4] += flats seen, counters[5] += flats fixed
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_link_flats(D8Args a, const int16_t* __restrict__ in, int16_t* __restrict__ out,
             unsigned long long* counters) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    unsigned is_flat = 0, fixed = 0;
    if (x < a.nx && y < a.ny) {
        const int64_t row = (int64_t)y * a.pitch;
        int c = in[row + x];
        const int gy = y + a.row_offset;
        const bool not_edge = (x > 0 && x < a.nx - 1 && gy > 0 && gy < a.global_ny - 1);
        if (c < 0 && c != PIT_CODE && not_edge) {              // :426-428
            is_flat = 1;
            const int allowed = -c;                            // :505-521 (bit k set)
            int ready = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                int xx = wrap_x(x + c_dc[k], a.nx);            // :445-465
                int nb = in[row + (int64_t)c_dr[k] * a.pitch + xx];
                int opp = 1 << ((k + 4) & 7);                  // code_opps[k]
                bool valid = (nb > 0) && (nb != opp);          // :491-498
                if (valid && ((allowed >> k) & 1)) ready += (1 << k);   // :529-536
            }
            if (ready > 0) { c = c_resolve[ready]; fixed = 1; }   // :538-548
        }
        out[row + x] = (int16_t)c;
    }
    unsigned f = __ballot_sync(0xffffffffu, is_flat);
    unsigned g = __ballot_sync(0xffffffffu, fixed);
    if (((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0) {
        if (f) atomicAdd(&counters[4], (unsigned long long)__popc(f));
        if (g) atomicAdd(&counters[5], (unsigned long long)__popc(g));
    }
}

// ---------------------------------------------------------------------------
// uint8 codes, parent IDs, ds, dw   (one pass)
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint8_t valid_code(int c) {
    // d8_global.py:141-144: out of range -> 0, then valid_code_map
    if (c < 0 || c > 128) return 0;
    return (c != 0 && (c & (c - 1)) == 0) ? (uint8_t)c : (uint8_t)0;
}

__device__ __forceinline__ float len_by_code(int code, float diag, float ew, float ns, float none) {
    if (code & 0x55) return diag;     // 1, 4, 16, 64
    if (code & 0x22) return ew;       // 2, 32
    if (code & 0x88) return ns;       // 8, 128
    return none;
}

__global__ void __launch_bounds__(256)
k_finalize(D8Args a, const int16_t* __restrict__ raw, const uint8_t* code_in,
           uint8_t* code, int32_t* __restrict__ parent,
           float* __restrict__ ds, float* __restrict__ dw,
           const float* __restrict__ dx, const float* __restrict__ dy,
           const float* __restrict__ dd) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= a.nx || y >= a.ny) return;
    const int64_t i = (int64_t)y * a.pitch + x;
    const int c = raw ? (int)valid_code(raw[i]) : (int)code_in[i];
    code[i] = (uint8_t)c;
    // update_parent_ID_grid, d8_global.py:574-644 (int32 calendar IDs)
    const int gy = y + a.row_offset;
    const int nx = a.nx;
    const long long n_pixels = (long long)nx * a.global_ny;
    long long pid = 0;
    if (c > 0) {
        const int k = __ffs(c) - 1;
        const int incs[8] = {-nx + 1, 1, nx + 1, nx, nx - 1, -1, -nx - 1, -nx};
        pid = (long long)gy * nx + x + incs[k];
        if (x == nx - 1) {
            if (a.lr_periodic) { if (c == 1 || c == 2 || c == 4) pid -= nx; }
            else pid = 0;
        }
        if (x == 0) {
            if (a.lr_periodic) { if (c == 16 || c == 32 || c == 64) pid += nx; }
            else pid = 0;
        }
        if (gy == 0) {
            if (a.tb_periodic) { if (c == 1 || c == 64 || c == 128) pid += n_pixels; }
            else pid = 0;
        }
        if (gy == a.global_ny - 1) {
            if (a.tb_periodic) { if (c == 4 || c == 8 || c == 16) pid -= n_pixels; }
            else pid = 0;
        }
    }
    parent[i] = (int32_t)pid;      // exact while the global grid has < 2^31 cells
    // update_flow_width_grid :863-935, update_flow_length_grid :949-1004
    const float rx = dx[y], ry = dy[y], rd = dd[y];
    dw[i] = len_by_code(c, rd, ry, rx, rx);
    ds[i] = len_by_code(c, rd, rx, ry, rx);
}

// bit j of inmask: the neighbour opposite to direction j holds code (1 << j),
// i.e. it drains into this cell and is the j-th term of the reference's
// k = 1..8 scatter (channels_base.py:1346-1361).
__global__ void __launch_bounds__(256)
k_inmask(D8Args a, const uint8_t* __restrict__ code, uint8_t* __restrict__ inmask) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= a.nx || y >= a.ny) return;
    const int64_t row = (int64_t)y * a.pitch;
    int m = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        int xx = x - c_dc[j];
        if (xx < 0 || xx >= a.nx) {
            if (!a.lr_periodic) continue;
            xx = wrap_x(xx, a.nx);
        }
        int c = code[row - (int64_t)c_dr[j] * a.pitch + xx];   // aprons: neighbour strip / wrap / 0
        if (c == (1 << j)) m |= (1 << j);
    }
    inmask[row + x] = (uint8_t)m;
}

// ---------------------------------------------------------------------------
// contributing area (update_area_grid, d8_global.py:1018-1261)
//
// The reference sweeps the whole grid once per level of the flow tree.  A
// cell's value is written exactly once, after all of its children, as
//   A = float32(pixel_area); A += A_child[k] for k = 0..7 (UR,R,LR,B,LL,L,UL,T)
// so it can be evaluated in any topological order.  Each cell carries a
// counter of unfinished children; the thread that finishes a cell's last
// child computes the cell and keeps climbing.  No atomics touch the float32
// sums, only the integer counters, so the result is bit-identical.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_area_init(D8Args a, const uint8_t* __restrict__ code, const uint8_t* __restrict__ inmask,
            int32_t* __restrict__ pending, float* __restrict__ area) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= a.nx || y >= a.ny) return;
    const int64_t i = (int64_t)y * a.pitch + x;
    pending[i] = code[i] ? __popc((unsigned)inmask[i]) : -1;
    area[i] = 0.0f;
}

__device__ __forceinline__ float area_of(const D8Args& a, int x, int y, unsigned m,
                                         const float* area, float pa) {
    float acc = pa;                                           // :1209-1212
    // neighbour positions k = 0..7 <-> inmask bit (k + 4) & 7   (:1214-1244)
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int j = (k + 4) & 7;
        if ((m >> j) & 1u) {
            int xx = wrap_x(x + c_dc[k], a.nx);
            int yy = y + c_dr[k];                             // "%" wrap, :1105-1108
            if (yy < 0) yy += a.ny; else if (yy >= a.ny) yy -= a.ny;
            acc += __ldcg(&area[(int64_t)yy * a.pitch + xx]);
        }
    }
    return acc;
}

__global__ void __launch_bounds__(256)
k_area_run(D8Args a, const uint8_t* __restrict__ code, const uint8_t* __restrict__ inmask,
           int32_t* pending, float* area, const double* __restrict__ da_rows,
           double unit_div) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= a.nx || y >= a.ny) return;
    int64_t i = (int64_t)y * a.pitch + x;
    int c = code[i];
    if (c == 0 || inmask[i] != 0) return;          // only leaves start a climb
    unsigned m = 0;
    while (true) {
        const float pa = (float)(da_rows[y] / unit_div);      // :1064-1067
        area[i] = area_of(a, x, y, m, area, pa);
        // step to the parent (periodic % as in :1105-1108); single strip only
        const int k = __ffs(c) - 1;
        int px = wrap_x(x + c_dc[k], a.nx);
        int py = y + c_dr[k];
        if (py < 0) py += a.ny; else if (py >= a.ny) py -= a.ny;
        const int64_t pi = (int64_t)py * a.pitch + px;
        const int pc = code[pi];
        if (pc == 0) return;                        // parent never becomes "unknown"
        __threadfence();                            // publish area[i] before the count drops
        const int old = atomicSub(&pending[pi], 1);
        if (old != 1) return;                       // someone else will finish the parent
        __threadfence();
        x = px; y = py; i = pi; c = pc;
        m = inmask[i];
    }
}

// ---------------------------------------------------------------------------
// The same on a strip with neighbours: a flow path may cross strip boundaries
// any number of times, so the climb stops where the parent lies in a
// neighbour's rows, and resumes there after the next exchange of the edge rows
// of `area`.  A cell still sums its own value once, after all its children,
// local or remote, in the reference's order: bit-identical to one strip.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float area_of_strip(const D8Args& a, int x, int y, unsigned m,
                                               const float* area, float pa) {
    float acc = pa;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int j = (k + 4) & 7;
        if ((m >> j) & 1u)                                   // rows -1 and ny are the neighbours' edge rows
            acc += __ldcg(&area[(int64_t)(y + c_dr[k]) * a.pitch + (x + c_dc[k])]);
    }
    return acc;
}

// compute cell (x, y), whose children are all final, and keep climbing while this
// thread finishes the last child of the next parent and that parent is ours
__device__ void area_climb(const D8Args& a, const uint8_t* __restrict__ code, const uint8_t* __restrict__ inmask,
                           int32_t* pending, float* area, const double* __restrict__ da_rows, double unit_div,
                           int x, int y) {
    int64_t i = (int64_t)y * a.pitch + x;
    int c = code[i];
    unsigned m = inmask[i];
    while (true) {
        const float pa = (float)(da_rows[y] / unit_div);
        area[i] = area_of_strip(a, x, y, m, area, pa);
        const int k = __ffs(c) - 1;
        const int px = x + c_dc[k], py = y + c_dr[k];
        __threadfence();
        if (py < 0 || py >= a.ny) return;                    // the neighbour strip goes on after the exchange
        const int64_t pi = (int64_t)py * a.pitch + px;
        const int pc = code[pi];
        if (pc == 0) return;
        const int old = atomicSub(&pending[pi], 1);
        if (old != 1) return;
        __threadfence();
        x = px; y = py; i = pi; c = pc;
        m = inmask[i];
    }
}

__global__ void __launch_bounds__(256)
k_area_run_strip(D8Args a, const uint8_t* __restrict__ code, const uint8_t* __restrict__ inmask,
                 int32_t* pending, float* area, const double* __restrict__ da_rows, double unit_div) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= a.nx || y >= a.ny) return;
    const int64_t i = (int64_t)y * a.pitch + x;
    if (code[i] == 0 || inmask[i] != 0) return;              // leaves start
    area_climb(a, code, inmask, pending, area, da_rows, unit_div, x, y);
}

// After an exchange: every cell of the two edge rows looks at its children in the
// neighbour's edge row; a child that has a value now and was not counted before
// is counted (credited[side * pitch + x], one bit per direction), and the cell
// that thereby loses its last pending child is computed and climbs on.
__global__ void __launch_bounds__(256)
k_area_resume(D8Args a, const uint8_t* __restrict__ code, const uint8_t* __restrict__ inmask,
              int32_t* pending, float* area, const double* __restrict__ da_rows, double unit_div,
              uint8_t* credited, unsigned long long* counters) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int side = blockIdx.y;                             // 0: row 0 (children in row -1), 1: row ny-1
    if (x >= a.nx) return;
    if ((side == 0 && !a.has_north) || (side == 1 && !a.has_south)) return;
    const int y = side == 0 ? 0 : a.ny - 1;
    // (a one-row strip: both sides look at the same cell, each at its own remote children)
    const int64_t i = (int64_t)y * a.pitch + x;
    if (code[i] == 0) return;
    const unsigned m = inmask[i];
    unsigned cr = credited[side * a.pitch + x];
    unsigned n_new = 0;
    bool last = false;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int j = (k + 4) & 7;
        const int yy = y + c_dr[k];
        const bool remote = (side == 0) ? (yy == -1) : (yy == a.ny);
        if (!remote || !((m >> j) & 1u) || ((cr >> j) & 1u)) continue;
        if (__ldcg(&area[(int64_t)yy * a.pitch + (x + c_dc[k])]) > 0.0f) {
            cr |= 1u << j;
            ++n_new;
            if (atomicSub(&pending[i], 1) == 1) last = true;
        }
    }
    if (n_new) {
        credited[side * a.pitch + x] = (uint8_t)cr;
        atomicAdd(&counters[11], (unsigned long long)n_new);
    }
    if (last) {
        __threadfence();
        area_climb(a, code, inmask, pending, area, da_rows, unit_div, x, y);
    }
}

__global__ void __launch_bounds__(256)
k_area_count_unresolved(D8Args a, const uint8_t* __restrict__ code,
                        const float* __restrict__ area, unsigned long long* counters) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    unsigned bad = 0;
    if (x < a.nx && y < a.ny) {
        const int64_t i = (int64_t)y * a.pitch + x;
        bad = (code[i] != 0 && area[i] == 0.0f);
    }
    unsigned b = __ballot_sync(0xffffffffu, bad);
    if (((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0 && b)
        atomicAdd(&counters[8], (unsigned long long)__popc(b));
}

// copy one owned row of a raster to an apron row (single-strip wrap)
template <typename T>
__global__ void k_copy_row(T* dst, const T* src, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static D8Args d8_args(tf_handle h) {
    D8Args a;
    a.nx = h->g.nx; a.ny = h->g.ny; a.pitch = h->g.pitch;
    a.row_offset = h->g.row_offset; a.global_ny = h->g.global_ny;
    a.has_north = h->g.has_north; a.has_south = h->g.has_south;
    a.lr_periodic = h->lr_periodic; a.tb_periodic = h->tb_periodic;
    return a;
}

static dim3 d8_block() { return dim3(64, 4); }
static dim3 d8_grid(tf_handle h) {
    return dim3((h->g.nx + 63) / 64, (h->g.ny + 3) / 4);
}

template <typename T>
static int fill_wrap_aprons(tf_handle h, Raster<T>& r) {
    // np.roll / "%" semantics: row -1 is the last row, row ny the first.
    // A strip with a neighbour keeps whatever the exchange put there.
    const Geom& g = h->g;
    const int nb = (g.pitch + 255) / 256;
    if (!g.has_north)
        k_copy_row<T><<<nb, 256, 0, h->stream>>>(r.p - g.pitch, r.p + (int64_t)(g.ny - 1) * g.pitch, g.pitch);
    if (!g.has_south)
        k_copy_row<T><<<nb, 256, 0, h->stream>>>(r.p + (int64_t)g.ny * g.pitch, r.p, g.pitch);
    TF_CUDA(cudaGetLastError());
    return 0;
}

template <typename T>
static int zero_aprons(tf_handle h, Raster<T>& r) {
    const Geom& g = h->g;
    if (!g.has_north)
        TF_CUDA(cudaMemsetAsync(r.p - g.pitch, 0, sizeof(T) * g.pitch, h->stream));
    if (!g.has_south)
        TF_CUDA(cudaMemsetAsync(r.p + (int64_t)g.ny * g.pitch, 0, sizeof(T) * g.pitch, h->stream));
    return 0;
}

static int upload_rows(tf_handle h, float** dev, const float* host) {
    if (!*dev) TF_CUDA(cudaMalloc((void**)dev, sizeof(float) * h->g.ny));
    TF_CUDA(cudaMemcpyAsync(*dev, host, sizeof(float) * h->g.ny, cudaMemcpyHostToDevice, h->stream));
    return 0;
}

static int d8_alloc(tf_handle h) {
    TF_TRY(raster_alloc(h, h->code, 0));
    TF_TRY(raster_alloc(h, h->inmask, 0));
    TF_TRY(raster_alloc(h, h->parent, 0));
    TF_TRY(raster_alloc(h, h->ds0, 0));
    TF_TRY(raster_alloc(h, h->dw, 0));
    return 0;
}

// codes (raw int16 or ready uint8) -> code, parent, ds, dw, inmask
static int d8_derive(tf_handle h, bool from_raw) {
    D8Args a = d8_args(h);
    k_finalize<<<d8_grid(h), d8_block(), 0, h->stream>>>(
        a, from_raw ? h->raw.p : nullptr, h->code.p, h->code.p, h->parent.p,
        h->ds0.p, h->dw.p, h->dx, h->dy, h->dd);
    TF_CUDA(cudaGetLastError());
    // aprons of the code grid: wrapped rows when periodic top-bottom, else 0
    if (h->tb_periodic) TF_TRY(fill_wrap_aprons(h, h->code));
    else TF_TRY(zero_aprons(h, h->code));
    if (h->nccl_comm) {
        void* arr[1] = {h->code.p};
        size_t es[1] = {1};
        TF_TRY(comm_exchange_rows(h, h->stream, arr, es, 1));
    }
    k_inmask<<<d8_grid(h), d8_block(), 0, h->stream>>>(a, h->code.p, h->inmask.p);
    TF_CUDA(cudaGetLastError());
    TF_CUDA(cudaStreamSynchronize(h->stream));
    h->have_d8 = true;
    h->have_area = false;
    return 0;
}

extern "C" int tf_d8_set_halo_dem(tf_handle h, const float* north_row, const float* south_row) {
    TF_TRY(use_device(h));
    TF_TRY(raster_alloc(h, h->dem, 0));
    const Geom& g = h->g;
    if (north_row)
        TF_CUDA(cudaMemcpyAsync(h->dem.p - g.pitch, north_row, sizeof(float) * g.nx, cudaMemcpyHostToDevice, h->stream));
    if (south_row)
        TF_CUDA(cudaMemcpyAsync(h->dem.p + (int64_t)g.ny * g.pitch, south_row, sizeof(float) * g.nx, cudaMemcpyHostToDevice, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int tf_d8_set_row0_sizes(tf_handle h, float dx0, float dy0, float dd0) {
    TF_CHECK(h != nullptr, "null handle");
    h->row0_len[0] = dx0; h->row0_len[1] = dy0; h->row0_len[2] = dd0;
    h->have_row0_len = true;
    return 0;
}

extern "C" int tf_d8_build(tf_handle h, const float* dem, const float* dx, const float* dy,
                           const float* dd, float nodata, int link_flats, int lr_periodic,
                           int tb_periodic, int64_t* stats_out) {
    TF_TRY(use_device(h));
    TF_CHECK(dem && dx && dy && dd, "tf_d8_build: NULL input");
    const Geom& g = h->g;
    h->lr_periodic = lr_periodic != 0;
    h->tb_periodic = tb_periodic != 0;
    uint8_t resolve[256];
    make_resolve(resolve);
    TF_CUDA(cudaMemcpyToSymbolAsync(c_resolve, resolve, 256, 0, cudaMemcpyHostToDevice, h->stream));
    TF_TRY(raster_alloc(h, h->dem, 0));
    TF_TRY(raster_alloc(h, h->raw, 0));
    TF_TRY(raster_alloc(h, h->raw2, 0));
    TF_TRY(d8_alloc(h));
    TF_TRY(raster_upload(h, h->dem, dem));
    TF_TRY(upload_rows(h, &h->dx, dx));
    TF_TRY(upload_rows(h, &h->dy, dy));
    TF_TRY(upload_rows(h, &h->dd, dd));
    TF_TRY(fill_wrap_aprons(h, h->dem));
    if (h->nccl_comm) {
        void* arr[1] = {h->dem.p};
        size_t es[1] = {sizeof(float)};
        TF_TRY(comm_exchange_rows(h, h->stream, arr, es, 1));
    }
    TF_CUDA(cudaMemsetAsync(h->counters, 0, 16 * sizeof(unsigned long long), h->stream));
    D8Args a = d8_args(h);
    // d8_global.py:209-211: the distances of row 0 -- of the GLOBAL grid -- are
    // used for every row.  A strip further south learns them from
    // tf_d8_set_row0_sizes or, failing that, from rank 0 over the communicator.
    float len0[3] = {dx[0], dy[0], dd[0]};
    if (h->have_row0_len) memcpy(len0, h->row0_len, sizeof(len0));
    else if (h->nccl_comm) {
        unsigned long long bits[3] = {0, 0, 0};
        if (h->rank == 0)
            for (int k = 0; k < 3; ++k) { uint32_t b; memcpy(&b, &len0[k], 4); bits[k] = b; }
        TF_CUDA(cudaMemcpyAsync(h->counters + 12, bits, sizeof(bits), cudaMemcpyHostToDevice, h->stream));
        TF_TRY(comm_allreduce_u64(h, h->stream, h->counters + 12, 3));   // sum with zeros = broadcast
        TF_CUDA(cudaMemcpyAsync(bits, h->counters + 12, sizeof(bits), cudaMemcpyDeviceToHost, h->stream));
        TF_CUDA(cudaStreamSynchronize(h->stream));
        for (int k = 0; k < 3; ++k) { const uint32_t b = (uint32_t)bits[k]; memcpy(&len0[k], &b, 4); }
    }
    k_start_codes<<<d8_grid(h), d8_block(), 0, h->stream>>>(
        a, h->dem.p, h->raw.p, len0[0], len0[1], len0[2], nodata, link_flats != 0, h->counters);
    TF_CUDA(cudaGetLastError());
    int64_t n_reps = 0;
    if (link_flats) {
        int16_t* cur = h->raw.p;
        int16_t* nxt = h->raw2.p;
        while (true) {
            ++n_reps;
            // aprons of the working codes: wrapped rows ("%" in the reference)
            {
                Raster<int16_t> tmp; tmp.p = cur;
                TF_TRY(fill_wrap_aprons(h, tmp));
                if (h->nccl_comm) {
                    void* arr[1] = {cur};
                    size_t es[1] = {sizeof(int16_t)};
                    TF_TRY(comm_exchange_rows(h, h->stream, arr, es, 1));
                }
            }
            TF_CUDA(cudaMemsetAsync(h->counters + 4, 0, 2 * sizeof(unsigned long long), h->stream));
            k_link_flats<<<d8_grid(h), d8_block(), 0, h->stream>>>(a, cur, nxt, h->counters);
            TF_CUDA(cudaGetLastError());
            TF_TRY(comm_allreduce_u64(h, h->stream, h->counters + 4, 2));   // all strips stop together
            unsigned long long c2[2];
            TF_CUDA(cudaMemcpyAsync(c2, h->counters + 4, sizeof(c2), cudaMemcpyDeviceToHost, h->stream));
            TF_CUDA(cudaStreamSynchronize(h->stream));
            const unsigned long long n_flats = c2[0], n_fix = c2[1];
            if (n_flats == 0) break;                           // :431-433
            if (n_fix) { int16_t* t = cur; cur = nxt; nxt = t; }
            if (n_fix == n_flats || n_fix == 0) break;         // :555-557
        }
        if (cur != h->raw.p) {
            TF_CUDA(cudaMemcpyAsync(h->raw.base, h->raw2.base, sizeof(int16_t) * (size_t)g.n_alloc,
                                    cudaMemcpyDeviceToDevice, h->stream));
        }
    }
    if (stats_out) {
        TF_TRY(comm_allreduce_u64(h, h->stream, h->counters, 3));
        unsigned long long c3[3];
        TF_CUDA(cudaMemcpyAsync(c3, h->counters, sizeof(c3), cudaMemcpyDeviceToHost, h->stream));
        TF_CUDA(cudaStreamSynchronize(h->stream));
        stats_out[0] = (int64_t)c3[0];
        stats_out[1] = (int64_t)c3[1];
        stats_out[2] = (int64_t)c3[2];
        stats_out[3] = n_reps;
    }
    return d8_derive(h, true);
}

extern "C" int tf_d8_set_codes(tf_handle h, const uint8_t* codes, const float* dx,
                               const float* dy, const float* dd, int lr_periodic,
                               int tb_periodic) {
    TF_TRY(use_device(h));
    TF_CHECK(codes && dx && dy && dd, "tf_d8_set_codes: NULL input");
    h->lr_periodic = lr_periodic != 0;
    h->tb_periodic = tb_periodic != 0;
    TF_TRY(d8_alloc(h));
    TF_TRY(raster_upload(h, h->code, codes));
    TF_TRY(upload_rows(h, &h->dx, dx));
    TF_TRY(upload_rows(h, &h->dy, dy));
    TF_TRY(upload_rows(h, &h->dd, dd));
    return d8_derive(h, false);
}

// update_area_grid on one strip that holds the whole grid (da on the device)
static int area_whole_grid(tf_handle h, const double* d_da, int units_km, int64_t* n_unresolved_out) {
    TF_TRY(raster_alloc(h, h->area, 0));
    TF_TRY(raster_alloc(h, h->pending, 0));
    D8Args a = d8_args(h);
    // the area sweep addresses neighbours with "%": aprons must wrap
    TF_TRY(fill_wrap_aprons(h, h->code));
    k_area_init<<<d8_grid(h), d8_block(), 0, h->stream>>>(a, h->code.p, h->inmask.p, h->pending.p, h->area.p);
    k_area_run<<<d8_grid(h), d8_block(), 0, h->stream>>>(a, h->code.p, h->inmask.p, h->pending.p,
                                                          h->area.p, d_da, units_km ? 1e6 : 1.0);
    TF_CUDA(cudaGetLastError());
    TF_CUDA(cudaMemsetAsync(h->counters + 8, 0, sizeof(unsigned long long), h->stream));
    k_area_count_unresolved<<<d8_grid(h), d8_block(), 0, h->stream>>>(a, h->code.p, h->area.p, h->counters);
    unsigned long long nbad = 0;
    TF_CUDA(cudaMemcpyAsync(&nbad, h->counters + 8, sizeof(nbad), cudaMemcpyDeviceToHost, h->stream));
    if (!h->tb_periodic) TF_TRY(zero_aprons(h, h->code));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    if (n_unresolved_out) *n_unresolved_out = (int64_t)nbad;
    h->have_area = true;
    return 0;
}

// Strips: flow paths cross strip boundaries and the float32 sums must see the
// true child values in the reference's order, so the area does not shard.
// Replicas instead: every rank gathers the uint8 code grid of all
// strips, accumulates the whole grid redundantly on its own GPU and keeps the
// rows it owns.
static int area_replicated(tf_handle h, const double* d_da, int units_km, int64_t* n_unresolved_out) {
    const Geom& g = h->g;
    TF_CHECK(h->nccl_comm != nullptr,
             "tf_d8_area: this strip has neighbours but no communicator (tf_comm_init)");
    TF_CHECK(h->nranks <= 64, "tf_d8_area: more than 64 strips");
    // who owns which rows
    unsigned long long tab[128];
    memset(tab, 0, sizeof(tab));
    tab[2 * h->rank] = (unsigned long long)g.row_offset;
    tab[2 * h->rank + 1] = (unsigned long long)g.ny;
    unsigned long long* d_tab = nullptr;
    TF_CUDA(cudaMalloc((void**)&d_tab, sizeof(tab)));
    TF_CUDA(cudaMemcpyAsync(d_tab, tab, sizeof(tab), cudaMemcpyHostToDevice, h->stream));
    TF_TRY(comm_allreduce_u64(h, h->stream, d_tab, 2 * h->nranks));
    TF_CUDA(cudaMemcpyAsync(tab, d_tab, sizeof(tab), cudaMemcpyDeviceToHost, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    cudaFree(d_tab);
    int first[64], rows[64];
    long long covered = 0;
    for (int r = 0; r < h->nranks; ++r) {
        first[r] = (int)tab[2 * r];
        rows[r] = (int)tab[2 * r + 1];
        TF_CHECK(first[r] == covered, "tf_d8_area: strips do not tile the grid (rank %d starts at row %d, "
                 "expected %lld)", r, first[r], covered);
        covered += rows[r];
    }
    TF_CHECK(covered == g.global_ny, "tf_d8_area: strips cover %lld of %d rows", covered, g.global_ny);
    // the whole grid on this GPU
    tf_handle T = nullptr;
    TF_TRY(tf_create(g.nx, g.global_ny, h->device, 0, &T));
    T->lr_periodic = h->lr_periodic;
    T->tb_periodic = h->tb_periodic;
    int rc = 0;
    double* d_da_all = nullptr;
    do {
        if ((rc = raster_alloc(T, T->code, 0))) break;
        if ((rc = raster_alloc(T, T->inmask, 0))) break;
        if (cudaMalloc((void**)&d_da_all, sizeof(double) * g.global_ny) != cudaSuccess) {
            tf_set_error("tf_d8_area: cudaMalloc failed"); rc = 1; break;
        }
        if (cudaStreamSynchronize(T->stream) != cudaSuccess) { tf_set_error("tf_d8_area: sync failed"); rc = 1; break; }
        if ((rc = comm_allgather_rows(h, h->stream, h->code.p, T->code.p, (size_t)g.pitch, first, rows))) break;
        if ((rc = comm_allgather_rows(h, h->stream, d_da, d_da_all, sizeof(double), first, rows))) break;
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { tf_set_error("tf_d8_area: gather failed"); rc = 1; break; }
        // aprons of the gathered code grid and its transpose, as d8_derive does for one strip
        if (T->tb_periodic) { if ((rc = fill_wrap_aprons(T, T->code))) break; }
        else if ((rc = zero_aprons(T, T->code))) break;
        k_inmask<<<d8_grid(T), d8_block(), 0, T->stream>>>(d8_args(T), T->code.p, T->inmask.p);
        T->have_d8 = true;
        if ((rc = area_whole_grid(T, d_da_all, units_km, n_unresolved_out))) break;
        if ((rc = raster_alloc(h, h->area, 0))) break;
        if (cudaMemcpyAsync(h->area.p, T->area.p + (int64_t)g.row_offset * g.pitch,
                            sizeof(float) * (size_t)g.ny * g.pitch, cudaMemcpyDeviceToDevice,
                            h->stream) != cudaSuccess ||
            cudaStreamSynchronize(h->stream) != cudaSuccess) {
            tf_set_error("tf_d8_area: copying the strip's rows failed"); rc = 1; break;
        }
        h->have_area = true;
    } while (0);
    if (d_da_all) cudaFree(d_da_all);
    std::string keep = tf_last_error();
    tf_destroy(T);
    if (rc) tf_set_error("%s", keep.c_str());
    return rc;
}

// Strips, owner computes: local climbs, then rounds of "exchange the edge rows of the
// area, resume" until no strip credits a new remote child.  The number of rounds is the
// largest number of strip boundaries one flow path crosses, plus one.
static int area_strips(tf_handle h, const double* d_da, int units_km, int64_t* n_unresolved_out) {
    const Geom& g = h->g;
    TF_CHECK(h->nccl_comm != nullptr,
             "tf_d8_area: this strip has neighbours but no communicator (tf_comm_init)");
    TF_TRY(raster_alloc(h, h->area, 0));
    TF_TRY(raster_alloc(h, h->pending, 0));
    D8Args a = d8_args(h);
    const double unit_div = units_km ? 1e6 : 1.0;
    uint8_t* credited = nullptr;
    TF_CUDA(cudaMalloc((void**)&credited, (size_t)2 * g.pitch));
    int rc = 0;
    long long rounds = 0;
    do {
        cudaMemsetAsync(credited, 0, (size_t)2 * g.pitch, h->stream);
        cudaMemsetAsync(h->area.base, 0, sizeof(float) * (size_t)g.n_alloc, h->stream);
        k_area_init<<<d8_grid(h), d8_block(), 0, h->stream>>>(a, h->code.p, h->inmask.p, h->pending.p, h->area.p);
        k_area_run_strip<<<d8_grid(h), d8_block(), 0, h->stream>>>(a, h->code.p, h->inmask.p, h->pending.p,
                                                                    h->area.p, d_da, unit_div);
        if (cudaGetLastError() != cudaSuccess) { tf_set_error("tf_d8_area: launch failed"); rc = 1; break; }
        while (true) {
            ++rounds;
            void* arr[1] = {h->area.p};
            size_t es[1] = {sizeof(float)};
            if ((rc = comm_exchange_rows(h, h->stream, arr, es, 1))) break;
            cudaMemsetAsync(h->counters + 11, 0, sizeof(unsigned long long), h->stream);
            k_area_resume<<<dim3((g.nx + 255) / 256, 2), 256, 0, h->stream>>>(a, h->code.p, h->inmask.p, h->pending.p,
                                                                              h->area.p, d_da, unit_div, credited,
                                                                              h->counters);
            if ((rc = comm_allreduce_u64(h, h->stream, h->counters + 11, 1))) break;
            unsigned long long n_new = 0;
            if (cudaMemcpyAsync(&n_new, h->counters + 11, sizeof(n_new), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
                cudaStreamSynchronize(h->stream) != cudaSuccess) {
                tf_set_error("tf_d8_area: %s", cudaGetErrorString(cudaGetLastError()));
                rc = 1;
                break;
            }
            if (n_new == 0) break;
        }
        if (rc) break;
        cudaMemsetAsync(h->counters + 8, 0, sizeof(unsigned long long), h->stream);
        k_area_count_unresolved<<<d8_grid(h), d8_block(), 0, h->stream>>>(a, h->code.p, h->area.p, h->counters);
        if ((rc = comm_allreduce_u64(h, h->stream, h->counters + 8, 1))) break;
        unsigned long long nbad = 0;
        cudaMemcpyAsync(&nbad, h->counters + 8, sizeof(nbad), cudaMemcpyDeviceToHost, h->stream);
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { tf_set_error("tf_d8_area: sync failed"); rc = 1; break; }
        if (n_unresolved_out) *n_unresolved_out = (int64_t)nbad;
        h->have_area = true;
        h->area_rounds = rounds;
    } while (0);
    cudaFree(credited);
    return rc;
}

extern "C" int tf_d8_area(tf_handle h, const double* da_rows, int units_km,
                          int64_t* n_unresolved_out) {
    TF_TRY(use_device(h));
    TF_CHECK(h->have_d8, "tf_d8_area: no D8 grid yet (call tf_d8_build first)");
    TF_CHECK(da_rows != nullptr, "tf_d8_area: da_rows is NULL");
    const Geom& g = h->g;
    double* d_da = nullptr;
    TF_CUDA(cudaMalloc((void**)&d_da, sizeof(double) * g.ny));
    TF_CUDA(cudaMemcpyAsync(d_da, da_rows, sizeof(double) * g.ny, cudaMemcpyHostToDevice, h->stream));
    // strips: owner computes (any size of the global grid); a periodic grid, whose flow paths wrap
    // around the outer edges, still takes the replicated whole-grid route (TF_B200_AREA_REPLICATED=1
    // forces it, for comparison)
    const char* e = getenv("TF_B200_AREA_REPLICATED");
    const bool strips = (g.has_north || g.has_south);
    const bool replicate = strips && (h->lr_periodic || h->tb_periodic || (e && e[0] == '1'));
    const int rc = !strips ? area_whole_grid(h, d_da, units_km, n_unresolved_out)
                           : (replicate ? area_replicated(h, d_da, units_km, n_unresolved_out)
                                        : area_strips(h, d_da, units_km, n_unresolved_out));
    cudaFree(d_da);
    return rc;
}

extern "C" long long tf_d8_area_rounds(tf_handle h) { return h ? h->area_rounds : -1; }

extern "C" int tf_d8_get(tf_handle h, int which, void* host_out) {
    TF_TRY(use_device(h));
    TF_CHECK(host_out != nullptr, "tf_d8_get: host_out is NULL");
    TF_CHECK(h->have_d8, "tf_d8_get: no D8 grid yet");
    switch (which) {
        case TF_D8_CODE: return raster_download<uint8_t>(h, h->code.p, (uint8_t*)host_out);
        case TF_D8_PARENT_ID:
            TF_CHECK((int64_t)h->g.nx * h->g.global_ny < ((int64_t)1 << 31),
                     "tf_d8_get: the global grid has >= 2^31 cells; the reference's int32 parent IDs "
                     "(d8_base.py:675) cannot represent it");
            return raster_download<int32_t>(h, h->parent.p, (int32_t*)host_out);
        case TF_D8_DS: return raster_download<float>(h, h->ds0.p, (float*)host_out);
        case TF_D8_DW: return raster_download<float>(h, h->dw.p, (float*)host_out);
        case TF_D8_AREA:
            TF_CHECK(h->have_area, "tf_d8_get: area not computed (call tf_d8_area)");
            return raster_download<float>(h, h->area.p, (float*)host_out);
        case TF_D8_INMASK: return raster_download<uint8_t>(h, h->inmask.p, (uint8_t*)host_out);
        case TF_D8_RAW:
            TF_CHECK(h->raw.p != nullptr, "tf_d8_get: raw codes exist only after tf_d8_build");
            return raster_download<int16_t>(h, h->raw.p, (int16_t*)host_out);
        default: break;
    }
    tf_set_error("tf_d8_get: unknown field %d", which);
    return 1;
}

int d8_free(tf_handle h) {
    raster_free(h->dem); raster_free(h->raw); raster_free(h->raw2);
    raster_free(h->code); raster_free(h->inmask); raster_free(h->parent);
    raster_free(h->ds0); raster_free(h->dw); raster_free(h->area);
    raster_free(h->pending);
    if (h->dx) cudaFree(h->dx);
    if (h->dy) cudaFree(h->dy);
    if (h->dd) cudaFree(h->dd);
    h->dx = h->dy = h->dd = nullptr;
    return 0;
}
