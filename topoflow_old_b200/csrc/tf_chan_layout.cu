// tf_chan_layout.cu -- turn the parameter grids into the device layout.
//
// Two layouts exist, chosen once in tf_chan_init:
//   generic : width, tan, cos, S_bed, roughness, ds as float64 rasters;
//   packed  : width, S_bed, roughness, ds as float32 rasters and the bank
//             angle as a one-byte index into a (tan, cos) table.
// The packed layout is used only when it is EXACT: every value must survive
// the float64 -> float32 -> float64 round trip (true for anything the
// reference read from an RTG file, which is float32 on disk,
// channels_base.py:2150-2153) and the grid must hold at most 255 distinct
// bank angles.  Arithmetic is float64 in both layouts; only the bytes read
// per cell differ.
//
// Everything here runs on the device: tf_chan_set_param copies a grid straight
// into a float64 staging raster, and the exactness scan, the table of distinct
// angles, the float32 packing, the per-row flow lengths and the conveyance
// factor are kernels over those rasters.  (Round 1 did all of it in host loops:
// 25 s at 16384^2.)
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <utility>

#include "tf_chan_host.cuh"

#define HASH_SLOTS 1024
#define HASH_EMPTY 0xffffffffffffffffull

static dim3 lay_block() { return dim3(64, 4); }
static dim3 lay_grid(const Geom& g) { return dim3((g.nx + 63) / 64, (g.ny + 3) / 4); }

// ---------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------
// flags[0] |= 1 when some owned cell of v does not survive float64 -> float32 -> float64
static __global__ void k_f32_exact(int nx, int ny, int pitch, const double* __restrict__ v, unsigned* flags) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    bool bad = false;
    if (x < nx && y < ny) {
        const double d = v[(int64_t)y * pitch + x];
        bad = !((double)(float)d == d);
    }
    if (__any_sync(0xffffffffu, bad) && ((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0) atomicOr(flags, 1u);
}

__device__ __forceinline__ unsigned hash_u64(unsigned long long k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33;
    return (unsigned)k & (HASH_SLOTS - 1);
}

// distinct bit patterns of `key` (and, for each, one `val` seen with it) into an
// open-addressing table; flags[1] |= 1 when more than `max_keys` turn up
static __global__ void k_distinct_insert(int nx, int ny, int pitch, const double* __restrict__ key,
                                         const double* __restrict__ val, unsigned long long* keys,
                                         unsigned long long* vals, unsigned* count, unsigned max_keys,
                                         unsigned* flags) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= nx || y >= ny) return;
    const int64_t i = (int64_t)y * pitch + x;
    const unsigned long long k = (unsigned long long)__double_as_longlong(key[i]);
    unsigned s = hash_u64(k);
    for (int probe = 0; probe < HASH_SLOTS; ++probe, s = (s + 1) & (HASH_SLOTS - 1)) {
        unsigned long long cur = keys[s];
        if (cur == k) return;
        if (cur == HASH_EMPTY) {
            if (*count > max_keys) { atomicOr(flags + 1, 1u); return; }
            cur = atomicCAS(keys + s, HASH_EMPTY, k);
            if (cur == HASH_EMPTY) {
                if (val) vals[s] = (unsigned long long)__double_as_longlong(val[i]);
                if (atomicAdd(count, 1u) + 1 > max_keys) atomicOr(flags + 1, 1u);
                return;
            }
            if (cur == k) return;
        }
    }
    atomicOr(flags + 1, 1u);
}

// aidx = slot_index[slot of the cell's key]; flags[2] |= 1 when `val` differs from the table's
static __global__ void k_distinct_index(int nx, int ny, int pitch, const double* __restrict__ key,
                                        const double* __restrict__ val, const unsigned long long* __restrict__ keys,
                                        const unsigned long long* __restrict__ vals,
                                        const uint8_t* __restrict__ slot_index, uint8_t* aidx, unsigned* flags) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= nx || y >= ny) return;
    const int64_t i = (int64_t)y * pitch + x;
    const unsigned long long k = (unsigned long long)__double_as_longlong(key[i]);
    unsigned s = hash_u64(k);
    for (int probe = 0; probe < HASH_SLOTS; ++probe, s = (s + 1) & (HASH_SLOTS - 1)) {
        if (keys[s] == k) {
            aidx[i] = slot_index[s];
            if (val && vals[s] != (unsigned long long)__double_as_longlong(val[i])) atomicOr(flags + 2, 1u);
            return;
        }
    }
    atomicOr(flags + 2, 1u);
}

// dst (whole raster, aprons and padding included) = pad; owned cells = (T)src or the scalar
template <typename T>
static __global__ void k_pack(int nx, int ny, int pitch, int64_t n_alloc, int64_t origin, const double* src_base,
                              double scalar, T pad, T* dst_base) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_alloc) return;
    const int64_t r = e - origin;
    T v = pad;
    if (r >= 0) {
        const int64_t y = r / pitch, x = r - y * pitch;
        if (y < ny && x < nx) v = (T)(src_base ? src_base[e] : scalar);
    }
    dst_base[e] = v;
}

// the flow width of padding cells and apron rows must never trip the overbank test
static __global__ void k_dw_padding(int nx, int ny, int pitch, int64_t n_alloc, int64_t origin, float* dw_base) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_alloc) return;
    const int64_t r = e - origin;
    bool owned = false;
    if (r >= 0) {
        const int64_t y = r / pitch, x = r - y * pitch;
        owned = (y < ny && x < nx);
    }
    if (!owned) dw_base[e] = 3.0e38f;
}

// Per row: are ds and dw functions of the D8 code class alone?  out[8 * y + c] = the value of class
// c (ds: diag, E/W, N/S, -; dw: diag, E/W, N/S, -), 0 when the class does not occur; flags[3] |= 1
// when a class takes two different values in one row.
static __global__ void __launch_bounds__(256)
k_len_rows(int nx, int pitch, const uint8_t* __restrict__ code, const double* __restrict__ ds_grid, double ds_scalar,
           const float* __restrict__ dw, float* out, unsigned* flags) {
    __shared__ float lo[6][256], hi[6][256];
    const int y = blockIdx.x;
    float mn[6], mx[6];
#pragma unroll
    for (int c = 0; c < 6; ++c) { mn[c] = INFINITY; mx[c] = -INFINITY; }
    for (int x = threadIdx.x; x < nx; x += blockDim.x) {
        const int64_t i = (int64_t)y * pitch + x;
        const int cd = code[i];
        const int ks = (cd & 0x55) ? 0 : ((cd & 0x88) ? 2 : 1);
        const int kw = (cd & 0x55) ? 3 : ((cd & 0x22) ? 4 : 5);
        const float vs = (float)(ds_grid ? ds_grid[i] : ds_scalar), vw = dw[i];
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            const bool hit = (c == ks) || (c == kw);
            const float v = (c < 3) ? vs : vw;
            if (hit) { mn[c] = fminf(mn[c], v); mx[c] = fmaxf(mx[c], v); }
        }
    }
#pragma unroll
    for (int c = 0; c < 6; ++c) { lo[c][threadIdx.x] = mn[c]; hi[c][threadIdx.x] = mx[c]; }
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                lo[c][threadIdx.x] = fminf(lo[c][threadIdx.x], lo[c][threadIdx.x + s]);
                hi[c][threadIdx.x] = fmaxf(hi[c][threadIdx.x], hi[c][threadIdx.x + s]);
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        bool bad = false;
        for (int c = 0; c < 6; ++c) {
            const bool present = lo[c][0] <= hi[c][0];
            if (present && lo[c][0] != hi[c][0]) bad = true;
            // NaN lengths never compare equal: not a function of the class
            out[8 * y + (c < 3 ? c : c + 1)] = present ? lo[c][0] : 0.0f;
        }
        out[8 * y + 3] = out[8 * y + 7] = 0.0f;
        if (bad) atomicOr(flags + 3, 1u);
    }
}

// kinematic + Manning: conveyance factor sqrt(S_bed) / n (both operations correctly rounded)
static __global__ void k_kfac(int nx, int ny, int pitch, int64_t n_alloc, int64_t origin, const double* S_base,
                              double S_scalar, const double* n_base, double n_scalar, double* out_base) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_alloc) return;
    const int64_t r = e - origin;
    double v = 1.0;
    if (r >= 0) {
        const int64_t y = r / pitch, x = r - y * pitch;
        if (y < ny && x < nx) v = sqrt(S_base ? S_base[e] : S_scalar) / (n_base ? n_base[e] : n_scalar);
    }
    out_base[e] = v;
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static inline const double* stage_base(const HostParam& p) { return p.stage.base; }   // NULL: scalar

static int all_f32_exact(tf_handle h, const HostParam* const* ps, int n, unsigned* d_flags, bool* out) {
    const Geom& g = h->g;
    bool ok = true;
    for (int k = 0; k < n; ++k) {
        if (ps[k]->stage.p)
            k_f32_exact<<<lay_grid(g), lay_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, ps[k]->stage.p, d_flags);
        else ok = ok && ((double)(float)ps[k]->scalar == ps[k]->scalar);
    }
    TF_CUDA(cudaGetLastError());
    unsigned f = 0;
    TF_CUDA(cudaMemcpyAsync(&f, d_flags, sizeof(f), cudaMemcpyDeviceToHost, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    *out = ok && !(f & 1u);
    return 0;
}

// Distinct values of a staged grid -> table slots; fills host copies of the slot keys / values.
static int distinct_of(tf_handle h, const double* key, const double* val, unsigned long long* d_keys,
                       unsigned long long* d_vals, unsigned* d_flags, std::vector<unsigned long long>& keys,
                       std::vector<unsigned long long>& vals, bool* too_many) {
    const Geom& g = h->g;
    TF_CUDA(cudaMemsetAsync(d_keys, 0xff, sizeof(unsigned long long) * HASH_SLOTS, h->stream));
    TF_CUDA(cudaMemsetAsync(d_vals, 0, sizeof(unsigned long long) * HASH_SLOTS, h->stream));
    TF_CUDA(cudaMemsetAsync(d_flags + 4, 0, sizeof(unsigned), h->stream));                     // the count
    k_distinct_insert<<<lay_grid(g), lay_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, key, val, d_keys, d_vals,
                                                                  d_flags + 4, 255u, d_flags);
    TF_LAUNCHED(h, h->stream, "k_distinct_insert");
    keys.resize(HASH_SLOTS); vals.resize(HASH_SLOTS);
    unsigned f[5];
    TF_CUDA(cudaMemcpyAsync(keys.data(), d_keys, sizeof(unsigned long long) * HASH_SLOTS, cudaMemcpyDeviceToHost, h->stream));
    TF_CUDA(cudaMemcpyAsync(vals.data(), d_vals, sizeof(unsigned long long) * HASH_SLOTS, cudaMemcpyDeviceToHost, h->stream));
    TF_CUDA(cudaMemcpyAsync(f, d_flags, sizeof(f), cudaMemcpyDeviceToHost, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    *too_many = (f[1] & 1u) != 0;
    return 0;
}

extern "C" int tf_chan_set_param(tf_handle h, int which, const double* grid, double scalar,
                                 int is_grid_in_reference) {
    TF_TRY(use_device(h));
    TF_CHECK(which >= 0 && which < TF_P_COUNT, "tf_chan_set_param: unknown parameter %d", which);
    HostParam& p = h->hp[which];
    if (grid) {
        TF_TRY(raster_alloc(h, p.stage, 0));
        TF_TRY(raster_upload(h, p.stage, grid));          // straight into the pitched raster, no host copy
    } else raster_free(p.stage);
    p.scalar = scalar;
    p.is_grid = is_grid_in_reference != 0;
    p.set = true;
    if (which == TF_P_TAN_ANGLE || which == TF_P_COS_ANGLE) h->angle_table_n = -1;   // tan / cos given directly
    h->chan_ready = false;
    return 0;
}

// The bank angle itself [rad], as the reference holds it (channels_base.py:2191): the library
// returns the distinct values (at most 255; *n_out = -1 when there are more) so that the caller
// evaluates tan and cos with its own libm for those few only, instead of over the whole grid,
// and hands them back with tf_chan_set_angle_table.
extern "C" int tf_chan_set_angle(tf_handle h, const double* grid, double scalar, int is_grid_in_reference,
                                 double* distinct_out, int* n_out) {
    TF_TRY(use_device(h));
    TF_CHECK(distinct_out && n_out, "tf_chan_set_angle: NULL output");
    const Geom& g = h->g;
    h->angle_is_grid = is_grid_in_reference != 0;
    h->angle_table_n = -1;
    h->chan_ready = false;
    *n_out = -1;
    if (!grid) {
        raster_free(h->aidx);
        h->angle_values.assign(1, scalar);
        h->angle_scalar = true;
        distinct_out[0] = scalar;
        *n_out = 1;
        return 0;
    }
    h->angle_scalar = false;
    Raster<double> st;
    TF_TRY(raster_alloc(h, st, 0));
    unsigned long long *d_keys = nullptr, *d_vals = nullptr;
    unsigned* d_flags = reinterpret_cast<unsigned*>(h->counters);
    int rc = 0;
    do {
        if ((rc = raster_upload(h, st, grid))) break;
        if (cudaMalloc((void**)&d_keys, sizeof(unsigned long long) * HASH_SLOTS * 2) != cudaSuccess) { tf_set_error("tf_chan_set_angle: cudaMalloc"); rc = 1; break; }
        d_vals = d_keys + HASH_SLOTS;
        cudaMemsetAsync(d_flags, 0, 8 * sizeof(unsigned), h->stream);
        std::vector<unsigned long long> keys, vals;
        bool too_many = false;
        if ((rc = distinct_of(h, st.p, nullptr, d_keys, d_vals, d_flags, keys, vals, &too_many))) break;
        if (too_many) break;                                    // *n_out stays -1: use tan / cos grids
        // table order = ascending angle: the same table whatever order the cells were seen in
        std::vector<std::pair<double, int>> order;
        for (int s = 0; s < HASH_SLOTS; ++s)
            if (keys[s] != HASH_EMPTY) { double v; memcpy(&v, &keys[s], 8); order.push_back({v, s}); }
        std::sort(order.begin(), order.end(), [](const std::pair<double, int>& a, const std::pair<double, int>& b) {
            unsigned long long x, y; memcpy(&x, &a.first, 8); memcpy(&y, &b.first, 8);
            return (a.first < b.first) || (!(b.first < a.first) && x < y);
        });
        std::vector<uint8_t> slot_index(HASH_SLOTS, 0);
        h->angle_values.clear();
        for (size_t k = 0; k < order.size(); ++k) {
            slot_index[order[k].second] = (uint8_t)(k + 1);     // entry 0 stays the rectangle
            h->angle_values.push_back(order[k].first);
            distinct_out[k] = order[k].first;
        }
        uint8_t* d_si = nullptr;
        if (cudaMalloc((void**)&d_si, HASH_SLOTS) != cudaSuccess) { tf_set_error("tf_chan_set_angle: cudaMalloc"); rc = 1; break; }
        cudaMemcpyAsync(d_si, slot_index.data(), HASH_SLOTS, cudaMemcpyHostToDevice, h->stream);
        if ((rc = raster_alloc(h, h->aidx, 0))) { cudaFree(d_si); break; }
        cudaMemsetAsync(h->aidx.base, 0, (size_t)g.n_alloc, h->stream);
        k_distinct_index<<<lay_grid(g), lay_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, st.p, nullptr, d_keys, d_vals,
                                                                     d_si, h->aidx.p, d_flags);
        cudaStreamSynchronize(h->stream);
        cudaFree(d_si);
        if (cudaGetLastError() != cudaSuccess) { tf_set_error("tf_chan_set_angle: index kernel failed"); rc = 1; break; }
        *n_out = (int)order.size();
    } while (0);
    if (d_keys) cudaFree(d_keys);
    raster_free(st);
    return rc;
}

extern "C" int tf_chan_set_angle_table(tf_handle h, const double* tan_values, const double* cos_values, int n) {
    TF_CHECK(h != nullptr, "null handle");
    TF_CHECK(tan_values && cos_values, "tf_chan_set_angle_table: NULL table");
    TF_CHECK(n == (int)h->angle_values.size() && n >= 1 && n <= 255,
             "tf_chan_set_angle_table: %d values, tf_chan_set_angle found %d", n, (int)h->angle_values.size());
    h->angle_tan.assign(tan_values, tan_values + n);
    h->angle_cos.assign(cos_values, cos_values + n);
    h->angle_table_n = n;
    h->hp[TF_P_TAN_ANGLE].set = h->hp[TF_P_COS_ANGLE].set = true;
    h->hp[TF_P_TAN_ANGLE].is_grid = h->hp[TF_P_COS_ANGLE].is_grid = h->angle_is_grid;
    raster_free(h->hp[TF_P_TAN_ANGLE].stage); raster_free(h->hp[TF_P_COS_ANGLE].stage);
    return 0;
}

static __global__ void k_expand_lut(int nx, int ny, int pitch, const uint8_t* __restrict__ aidx,
                                    const double* __restrict__ lut, double* tan_out, double* cos_out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= nx || y >= ny) return;
    const int64_t i = (int64_t)y * pitch + x;
    tan_out[i] = lut[aidx[i]];
    cos_out[i] = lut[256 + aidx[i]];
}

template <typename T>
static int pack_param(tf_handle h, Raster<T>& dst, const HostParam& p, T pad) {
    const Geom& g = h->g;
    TF_TRY(raster_alloc(h, dst, 0));
    const unsigned nb = (unsigned)((g.n_alloc + 255) / 256);
    k_pack<T><<<nb, 256, 0, h->stream>>>(g.nx, g.ny, g.pitch, g.n_alloc, g.origin, stage_base(p), p.scalar, pad, dst.base);
    TF_LAUNCHED(h, h->stream, "k_pack");
    return 0;
}

int chan_layout_init(tf_handle h, int allow_packed) {
    const Geom& g = h->g;
    HostParam& W = h->hp[TF_P_WIDTH];
    HostParam& T = h->hp[TF_P_TAN_ANGLE];
    HostParam& C = h->hp[TF_P_COS_ANGLE];
    HostParam& S = h->hp[TF_P_SLOPE];
    HostParam& N = h->hp[TF_P_ROUGH];
    HostParam& D = h->hp[TF_P_DS];
    const bool angle_table = h->angle_table_n > 0;
    TF_CHECK(angle_table || ((T.stage.p == nullptr) == (C.stage.p == nullptr)),
             "tf_chan_init: tan and cos of the bank angle must both be grids or both scalars");
    unsigned* d_flags = reinterpret_cast<unsigned*>(h->counters);
    TF_CUDA(cudaMemsetAsync(d_flags, 0, 8 * sizeof(unsigned), h->stream));
    h->rcp_ok = true;
    bool packed = false;
    if (allow_packed) {
        const HostParam* ps[4] = {&W, &S, &N, &D};
        TF_TRY(all_f32_exact(h, ps, 4, d_flags, &packed));
    }
    std::vector<double> lut(768, 0.0);       // tan, cos, 1 / (2 tan); entry 0 = the rectangle
    lut[256] = 1.0;
    const unsigned nb_all = (unsigned)((g.n_alloc + 255) / 256);
    if (angle_table) {
        for (int k = 0; k < h->angle_table_n; ++k) { lut[k + 1] = h->angle_tan[k]; lut[256 + k + 1] = h->angle_cos[k]; }
        if (h->angle_scalar) {
            TF_TRY(raster_alloc(h, h->aidx, 0));
            k_pack<uint8_t><<<nb_all, 256, 0, h->stream>>>(g.nx, g.ny, g.pitch, g.n_alloc, g.origin, nullptr, 1.0,
                                                           (uint8_t)0, h->aidx.base);
            TF_LAUNCHED(h, h->stream, "k_pack");
        }
    } else if (packed) {
        if (!T.stage.p) {
            lut[1] = T.scalar; lut[257] = C.scalar;
            TF_TRY(raster_alloc(h, h->aidx, 0));
            k_pack<uint8_t><<<nb_all, 256, 0, h->stream>>>(g.nx, g.ny, g.pitch, g.n_alloc, g.origin, nullptr, 1.0,
                                                           (uint8_t)0, h->aidx.base);
            TF_LAUNCHED(h, h->stream, "k_pack");
        } else {
            // distinct (tan, cos) pairs: keyed on tan, cos checked to be a function of it
            unsigned long long* d_keys = nullptr;
            uint8_t* d_si = nullptr;
            TF_CUDA(cudaMalloc((void**)&d_keys, sizeof(unsigned long long) * HASH_SLOTS * 2));
            std::vector<unsigned long long> keys, vals;
            bool too_many = false;
            int rc = distinct_of(h, T.stage.p, C.stage.p, d_keys, d_keys + HASH_SLOTS, d_flags, keys, vals, &too_many);
            if (!rc && !too_many) {
                std::vector<std::pair<unsigned long long, int>> order;
                for (int s = 0; s < HASH_SLOTS; ++s) if (keys[s] != HASH_EMPTY) order.push_back({keys[s], s});
                std::sort(order.begin(), order.end());
                std::vector<uint8_t> slot_index(HASH_SLOTS, 0);
                for (size_t k = 0; k < order.size(); ++k) {
                    slot_index[order[k].second] = (uint8_t)(k + 1);
                    memcpy(&lut[k + 1], &keys[order[k].second], 8);
                    memcpy(&lut[256 + k + 1], &vals[order[k].second], 8);
                }
                if (cudaMalloc((void**)&d_si, HASH_SLOTS) != cudaSuccess) rc = 1;
                if (!rc) {
                    cudaMemcpyAsync(d_si, slot_index.data(), HASH_SLOTS, cudaMemcpyHostToDevice, h->stream);
                    rc = raster_alloc(h, h->aidx, 0);
                }
                if (!rc) {
                    cudaMemsetAsync(h->aidx.base, 0, (size_t)g.n_alloc, h->stream);
                    k_distinct_index<<<lay_grid(g), lay_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, T.stage.p, C.stage.p,
                                                                                 d_keys, d_keys + HASH_SLOTS, d_si,
                                                                                 h->aidx.p, d_flags);
                    unsigned f[4] = {0, 0, 0, 0};
                    cudaMemcpyAsync(f, d_flags, sizeof(f), cudaMemcpyDeviceToHost, h->stream);
                    if (cudaStreamSynchronize(h->stream) != cudaSuccess) rc = 1;
                    if (f[2] & 1u) packed = false;          // cos is not a function of tan here
                }
            } else if (!rc) packed = false;
            if (d_si) cudaFree(d_si);
            cudaFree(d_keys);
            if (rc) { tf_set_error("tf_chan_init: building the angle table failed"); return 1; }
        }
    }
    h->packed = packed ? 1 : 0;

    // third column: RN(1 / (2 tan)), the reciprocal chan_depth_rcp divides by; a denominator whose
    // reciprocal could overflow or underflow keeps the run off the reciprocal paths
    for (int k = 0; k < 256; ++k) {
        const double den = 2.0 * lut[k];
        if (den == 0.0) { lut[512 + k] = 0.0; continue; }
        if (!(fabs(den) > 1e-100 && fabs(den) < 1e100)) h->rcp_ok = false;
        lut[512 + k] = 1.0 / den;
    }
    if (h->lut) cudaFree(h->lut);
    h->lut = nullptr;
    TF_CUDA(cudaMalloc((void**)&h->lut, sizeof(double) * 768));
    TF_CUDA(cudaMemcpyAsync(h->lut, lut.data(), sizeof(double) * 768, cudaMemcpyHostToDevice, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));

    // padding cells get a harmless shape (unit rectangle, code 0 keeps them dry)
    if (packed) {
        TF_TRY(pack_param<float>(h, h->w32, W, 1.0f));
        TF_TRY(pack_param<float>(h, h->S32, S, 1.0f));
        TF_TRY(pack_param<float>(h, h->n32, N, 1.0f));
        TF_TRY(pack_param<float>(h, h->ds32, D, 1.0f));
    } else {
        TF_TRY(pack_param<double>(h, h->w64, W, 1.0));
        TF_TRY(pack_param<double>(h, h->S64, S, 1.0));
        TF_TRY(pack_param<double>(h, h->n64, N, 1.0));
        TF_TRY(pack_param<double>(h, h->ds64, D, 1.0));
        if (angle_table) {
            HostParam zero; zero.scalar = 0.0;
            HostParam one; one.scalar = 1.0;
            TF_TRY(pack_param<double>(h, h->tan64, zero, 0.0));
            TF_TRY(pack_param<double>(h, h->cos64, one, 1.0));
            k_expand_lut<<<lay_grid(g), lay_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, h->aidx.p, h->lut,
                                                                     h->tan64.p, h->cos64.p);
            TF_LAUNCHED(h, h->stream, "k_expand_lut");
        } else {
            TF_TRY(pack_param<double>(h, h->tan64, T, 0.0));
            TF_TRY(pack_param<double>(h, h->cos64, C, 1.0));
        }
    }
    // ds and dw as functions of the code and three lengths per row: use them only if that
    // reproduces every cell of both grids exactly
    if (h->len_rows) { cudaFree(h->len_rows); h->len_rows = nullptr; }
    if (packed && !getenv("TF_B200_NO_LEN_ROWS")) {
        float* d_rows = nullptr;
        TF_CUDA(cudaMalloc((void**)&d_rows, sizeof(float) * 8 * g.ny));
        k_len_rows<<<g.ny, 256, 0, h->stream>>>(g.nx, g.pitch, h->code.p, D.stage.p, D.scalar, h->dw.p, d_rows, d_flags);
        std::vector<float> r8((size_t)8 * g.ny);
        unsigned f[4] = {0, 0, 0, 0};
        cudaMemcpyAsync(r8.data(), d_rows, sizeof(float) * r8.size(), cudaMemcpyDeviceToHost, h->stream);
        cudaMemcpyAsync(f, d_flags, sizeof(f), cudaMemcpyDeviceToHost, h->stream);
        const cudaError_t e = cudaStreamSynchronize(h->stream);
        cudaFree(d_rows);
        TF_CHECK(e == cudaSuccess, "tf_chan_init: k_len_rows: %s", cudaGetErrorString(e));
        if (!(f[3] & 1u)) {
            std::vector<float> rows((size_t)TF_LEN_ROW * (g.ny + 2 * TF_APRON), 0.0f);
            for (int y = 0; y < g.ny; ++y) {
                float* r = rows.data() + (size_t)TF_LEN_ROW * (y + TF_APRON);
                memcpy(r, r8.data() + (size_t)8 * y, 8 * sizeof(float));
                double rcp[4] = {0.0, 0.0, 0.0, 0.0};              // floats 8..15: RN(1 / ds) as float64
                for (int k = 0; k < 3; ++k) {
                    const double v = (double)r[k];
                    if (v == 0.0) continue;                          // class absent in this row
                    if (!(v > 1e-30 && v < 1e30)) h->rcp_ok = false;
                    rcp[k] = 1.0 / v;
                }
                memcpy(r + 8, rcp, sizeof(rcp));
            }
            // apron slots: copies of the edge rows (a strip with neighbours receives the
            // neighbours' rows in chan_exchange_static)
            const size_t rb = TF_LEN_ROW * sizeof(float);
            for (int k = 0; k < TF_APRON; ++k) {
                memcpy(rows.data() + (size_t)TF_LEN_ROW * k, rows.data() + (size_t)TF_LEN_ROW * TF_APRON, rb);
                memcpy(rows.data() + (size_t)TF_LEN_ROW * (g.ny + TF_APRON + k),
                       rows.data() + (size_t)TF_LEN_ROW * (g.ny + TF_APRON - 1), rb);
            }
            TF_CUDA(cudaMalloc((void**)&h->len_rows, sizeof(float) * rows.size()));
            TF_CUDA(cudaMemcpy(h->len_rows, rows.data(), sizeof(float) * rows.size(), cudaMemcpyHostToDevice));
        }
    }
    k_dw_padding<<<nb_all, 256, 0, h->stream>>>(g.nx, g.ny, g.pitch, g.n_alloc, g.origin, h->dw.base);
    TF_LAUNCHED(h, h->stream, "k_dw_padding");
    // kinematic + Manning: conveyance factor sqrt(S_bed) / n, once
    raster_free(h->kfac);
    if (h->wave == TF_KINEMATIC && h->law == TF_MANNING && !getenv("TF_B200_NO_KFAC")) {
        TF_TRY(raster_alloc(h, h->kfac, 0));
        k_kfac<<<nb_all, 256, 0, h->stream>>>(g.nx, g.ny, g.pitch, g.n_alloc, g.origin, S.stage.base, S.scalar,
                                              N.stage.base, N.scalar, h->kfac.base);
        TF_LAUNCHED(h, h->stream, "k_kfac");
    }
    TF_CUDA(cudaStreamSynchronize(h->stream));
    // the staging rasters are done: the packed / generic rasters are the parameters from here on
    for (int k = 0; k < TF_P_COUNT; ++k) {
        raster_free(h->hp[k].stage);
        h->hp[k].set = false;
    }
    h->angle_table_n = -1;
    return 0;
}

// Strips: the fused marching step recomputes the volume and depth of the
// neighbour strip's edge row in this strip's apron row, so the apron row
// needs that row's static parameters as well (once, at initialisation).
int chan_exchange_static(tf_handle h) {
    if (!h->nccl_comm) return 0;
    const Geom& g = h->g;
    void* arr[8];
    size_t es[8];
    int n = 0;
    auto add = [&](void* p, size_t e) { if (p) { arr[n] = p; es[n] = e; ++n; } };
    add(h->w32.p, 4); add(h->S32.p, 4); add(h->n32.p, 4); add(h->aidx.p, 1);
    add(h->kfac.p, 8); add(h->inmask.p, 1);
    TF_TRY(comm_exchange_rows(h, h->stream, arr, es, n));
    float* L = h->len_rows;
    TF_TRY(comm_exchange_bytes(h, h->stream, L + TF_LEN_ROW * TF_APRON, L + TF_LEN_ROW * (TF_APRON - 1),
                               L + TF_LEN_ROW * (TF_APRON + g.ny - 1), L + TF_LEN_ROW * (TF_APRON + g.ny),
                               TF_LEN_ROW * sizeof(float)));
    double* A = h->da_rows;
    TF_TRY(comm_exchange_bytes(h, h->stream, A + TF_APRON, A + TF_APRON - 1,
                               A + TF_APRON + g.ny - 1, A + TF_APRON + g.ny, 8));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
}
