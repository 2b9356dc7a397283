// tf_chan_layout.cu -- turn the host parameter grids into the device layout.
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
#include <math.h>
#include <stdlib.h>

#include <map>
#include <utility>

#include "tf_chan_host.cuh"

static bool f32_exact(const HostParam& p) {
    if (p.grid.empty()) {
        const double v = p.scalar;
        return (double)(float)v == v;
    }
    for (double v : p.grid)
        if (!((double)(float)v == v)) return false;
    return true;
}

static inline double hp_at(const HostParam& p, size_t i) {
    return p.grid.empty() ? p.scalar : p.grid[i];
}

template <typename T>
static int fill_raster(tf_handle h, Raster<T>& r, const HostParam& p, T pad) {
    const Geom& g = h->g;
    TF_TRY(raster_alloc(h, r, 0));
    std::vector<T> host((size_t)g.n_alloc, pad);
    for (int y = 0; y < g.ny; ++y) {
        T* row = host.data() + g.origin + (size_t)y * g.pitch;
        for (int x = 0; x < g.nx; ++x) row[x] = (T)hp_at(p, (size_t)y * g.nx + x);
    }
    // on the handle's stream: raster_alloc has queued a memset there, and the
    // stream is non-blocking, so a copy on the default stream could overtake it
    TF_CUDA(cudaMemcpyAsync(r.base, host.data(), sizeof(T) * (size_t)g.n_alloc, cudaMemcpyHostToDevice, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
}

int chan_layout_init(tf_handle h, int allow_packed) {
    const Geom& g = h->g;
    const size_t n = (size_t)g.nx * g.ny;
    const HostParam& W = h->hp[TF_P_WIDTH];
    const HostParam& T = h->hp[TF_P_TAN_ANGLE];
    const HostParam& C = h->hp[TF_P_COS_ANGLE];
    const HostParam& S = h->hp[TF_P_SLOPE];
    const HostParam& N = h->hp[TF_P_ROUGH];
    const HostParam& D = h->hp[TF_P_DS];
    TF_CHECK(T.grid.empty() == C.grid.empty(),
             "tf_chan_init: tan and cos of the bank angle must both be grids or both scalars");

    h->rcp_ok = true;
    bool packed = allow_packed && f32_exact(W) && f32_exact(S) && f32_exact(N) && f32_exact(D);
    std::vector<uint8_t> aidx;
    std::vector<double> lut(768, 0.0);   // tan, cos, 1 / (2 tan)
    if (packed) {
        // entry 0 is the rectangle (tan 0, cos 1) the overbank rule switches to
        std::map<std::pair<uint64_t, uint64_t>, int> seen;
        auto bits = [](double v) { uint64_t b; memcpy(&b, &v, 8); return b; };
        lut[0] = 0.0; lut[256] = 1.0;
        seen[{bits(0.0), bits(1.0)}] = 0;
        int n_used = 1;
        const size_t n_ang = T.grid.empty() ? 1 : n;
        aidx.assign(n, 0);
        uint64_t last_t = ~0ull, last_c = ~0ull;
        int last_i = -1;
        for (size_t i = 0; i < n_ang && packed; ++i) {
            const uint64_t bt = bits(hp_at(T, i)), bc = bits(hp_at(C, i));
            int idx;
            if (bt == last_t && bc == last_c) idx = last_i;
            else {
                auto it = seen.find({bt, bc});
                if (it != seen.end()) idx = it->second;
                else if (n_used < 256) {
                    idx = n_used++;
                    seen[{bt, bc}] = idx;
                    lut[idx] = hp_at(T, i);
                    lut[256 + idx] = hp_at(C, i);
                } else { packed = false; break; }
                last_t = bt; last_c = bc; last_i = idx;
            }
            aidx[i] = (uint8_t)idx;
        }
        if (packed && T.grid.empty()) std::fill(aidx.begin(), aidx.end(), aidx[0]);
    }
    h->packed = packed ? 1 : 0;

    // padding cells get a harmless shape (unit rectangle, code 0 keeps them dry)
    if (packed) {
        TF_TRY(fill_raster<float>(h, h->w32, W, 1.0f));
        TF_TRY(fill_raster<float>(h, h->S32, S, 1.0f));
        TF_TRY(fill_raster<float>(h, h->n32, N, 1.0f));
        TF_TRY(fill_raster<float>(h, h->ds32, D, 1.0f));
        HostParam A;
        A.grid.assign(aidx.begin(), aidx.end());
        TF_TRY(fill_raster<uint8_t>(h, h->aidx, A, (uint8_t)0));
        // third column: RN(1 / (2 tan)), the reciprocal chan_depth_rcp divides by;
        // a denominator whose reciprocal could overflow or underflow keeps the
        // run on the tile kernels (ordinary division)
        h->rcp_ok = true;
        for (int k = 0; k < 256; ++k) {
            const double den = 2.0 * lut[k];
            if (den == 0.0) { lut[512 + k] = 0.0; continue; }
            if (!(fabs(den) > 1e-100 && fabs(den) < 1e100)) h->rcp_ok = false;
            lut[512 + k] = 1.0 / den;
        }
        if (h->lut) cudaFree(h->lut);
        TF_CUDA(cudaMalloc((void**)&h->lut, sizeof(double) * 768));
        TF_CUDA(cudaMemcpy(h->lut, lut.data(), sizeof(double) * 768, cudaMemcpyHostToDevice));
    } else {
        TF_TRY(fill_raster<double>(h, h->w64, W, 1.0));
        TF_TRY(fill_raster<double>(h, h->S64, S, 1.0));
        TF_TRY(fill_raster<double>(h, h->n64, N, 1.0));
        TF_TRY(fill_raster<double>(h, h->ds64, D, 1.0));
        TF_TRY(fill_raster<double>(h, h->tan64, T, 0.0));
        TF_TRY(fill_raster<double>(h, h->cos64, C, 1.0));
    }
    // the flow width of padding cells must never trip the overbank test
    {
        std::vector<float> dw((size_t)g.n_alloc);
        TF_CUDA(cudaMemcpy(dw.data(), h->dw.base, sizeof(float) * (size_t)g.n_alloc, cudaMemcpyDeviceToHost));
        // ds and dw as functions of the code and three lengths per row: use
        // them only if that reproduces every cell of both grids exactly
        if (h->len_rows) { cudaFree(h->len_rows); h->len_rows = nullptr; }
        if (packed && !getenv("TF_B200_NO_LEN_ROWS")) {
            std::vector<uint8_t> code((size_t)g.n_alloc);
            TF_CUDA(cudaMemcpy(code.data(), h->code.base, (size_t)g.n_alloc, cudaMemcpyDeviceToHost));
            std::vector<float> rows((size_t)TF_LEN_ROW * (g.ny + 2 * TF_APRON), 0.0f);
            bool ok = true;
            for (int y = 0; y < g.ny && ok; ++y) {
                const uint8_t* cr = code.data() + g.origin + (int64_t)y * g.pitch;
                const float* wr = dw.data() + g.origin + (int64_t)y * g.pitch;
                float* r = rows.data() + (size_t)TF_LEN_ROW * (y + TF_APRON);
                bool have[6] = {false, false, false, false, false, false};
                for (int x = 0; x < g.nx; ++x) {
                    const int cd = cr[x];
                    const int ks = (cd & 0x55) ? 0 : ((cd & 0x88) ? 2 : 1);        // ds class
                    const int kw = (cd & 0x55) ? 4 : ((cd & 0x22) ? 5 : 6);        // dw class
                    const float vs = (float)hp_at(D, (size_t)y * g.nx + x), vw = wr[x];
                    if (!have[ks]) { have[ks] = true; r[ks] = vs; }
                    else if (r[ks] != vs) { ok = false; break; }
                    if (!have[kw - 1]) { have[kw - 1] = true; r[kw] = vw; }
                    else if (r[kw] != vw) { ok = false; break; }
                }
            }
            if (ok) {
                // floats 8..15 of a row: RN(1 / ds) of the three classes as float64
                for (int y = 0; y < g.ny; ++y) {
                    float* r = rows.data() + (size_t)TF_LEN_ROW * (y + TF_APRON);
                    double rcp[4] = {0.0, 0.0, 0.0, 0.0};
                    for (int k = 0; k < 3; ++k) {
                        const double v = (double)r[k];
                        if (v == 0.0) continue;                  // class absent in this row
                        if (!(v > 1e-30 && v < 1e30)) h->rcp_ok = false;
                        rcp[k] = 1.0 / v;
                    }
                    memcpy(r + 8, rcp, sizeof(rcp));
                }
                // apron slots: copies of the edge rows (a strip with neighbours
                // receives the neighbours' rows in chan_exchange_static)
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
        for (int y = -TF_APRON; y < g.ny + TF_APRON; ++y) {
            float* row = dw.data() + g.origin + (int64_t)y * g.pitch;
            for (int x = 0; x < g.pitch; ++x)
                if (y < 0 || y >= g.ny || x >= g.nx) row[x] = 3.0e38f;
        }
        TF_CUDA(cudaMemcpy(h->dw.base, dw.data(), sizeof(float) * (size_t)g.n_alloc, cudaMemcpyHostToDevice));
    }
    // kinematic + Manning: conveyance factor sqrt(S_bed)/n, once, on the host
    raster_free(h->kfac);
    if (h->wave == TF_KINEMATIC && h->law == TF_MANNING && !getenv("TF_B200_NO_KFAC")) {
        HostParam K;
        K.grid.resize(n);
        for (size_t i = 0; i < n; ++i) K.grid[i] = sqrt(hp_at(S, i)) / hp_at(N, i);
        TF_TRY(fill_raster<double>(h, h->kfac, K, 1.0));
    }
    // free the host copies: the device rasters are the parameters from here on
    for (int k = 0; k < TF_P_COUNT; ++k) {
        h->hp[k].grid.clear();
        h->hp[k].grid.shrink_to_fit();
        h->hp[k].set = false;
    }
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
