// tf_fill.cu -- two neighbours of the routing path that run on the same D8
// structures:
//
//   tf_fill_pits      depression filling of a DEM.  The reference
//                     (utils/fill_pits.py:178-582) is the priority-flood of
//                     Wang & Liu with a binary heap: a serial algorithm.  Its
//                     result, however, is unique -- every cell ends at
//                     max(z, lowest pass height on any path to a start cell)
//                     -- and only ever copies elevations, so any evaluation
//                     order gives the same bits.  Here the surface is relaxed
//                     from above (Planchon & Darboux): W = +inf except on the
//                     start cells, then W <- max(z, min over the 8 neighbours
//                     of W) until nothing changes, many local iterations per
//                     tile in shared memory for every pass over the grid.
//
//   tf_satzone_dzw    the lateral-flow part of satzone_base.update_water_table
//                     (components/satzone_base.py:930-1008): dzw = dt (Rg -
//                     Q_gw / da), then dzw[p_k] += Q_gw[w_k] dt / da for k =
//                     1..8 -- the same scatter through the D8 index lists as
//                     update_flow_volume, done as a gather over the child mask
//                     in the same order.
#include <math.h>

#include "tf_common.cuh"

#define FT 32                 // tile edge
#define FILL_LOCAL_ITERS 64

// start cells (fill_pits.py:178-330): grid-edge cells with data, and interior
// cells with data next to a nodata cell; nodata / non-finite cells never change
__global__ void __launch_bounds__(256)
k_fill_init(int nx, int ny, int pitch, const float* __restrict__ z, float nodata, float* __restrict__ W) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= nx || y >= ny) return;
    const int64_t i = (int64_t)y * pitch + x;
    const float zc = z[i];
    const bool data = isfinite(zc) && zc > nodata;
    float w = INFINITY;
    if (!data) w = isfinite(zc) ? zc : nodata;          // closed from the start (:372-378); NaN -> nodata (:239-242)
    else if (x == 0 || y == 0 || x == nx - 1 || y == ny - 1) w = zc;
    else {
        float zs = INFINITY;
        for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx) {
                if (!dx && !dy) continue;
                float zn = z[i + (int64_t)dy * pitch + dx];
                if (!isfinite(zn)) zn = nodata;         // (:239-242: non-finite cells become nodata first)
                zs = fminf(zs, zn);
            }
        if (zs <= nodata) w = zc;                       // :318-321
    }
    W[i] = w;
}

__global__ void __launch_bounds__(FT * 8)
k_fill_relax(int nx, int ny, int pitch, const float* __restrict__ z, float nodata, float* W, int* changed) {
    __shared__ float sw[FT + 2][FT + 2];
    __shared__ float sz[FT][FT];
    __shared__ int s_changed, s_any;
    const int x0 = blockIdx.x * FT, y0 = blockIdx.y * FT;
    const int tid = threadIdx.y * FT + threadIdx.x;
    for (int k = tid; k < (FT + 2) * (FT + 2); k += FT * 8) {
        const int ly = k / (FT + 2), lx = k % (FT + 2);
        const int x = x0 + lx - 1, y = y0 + ly - 1;
        float w = INFINITY;                              // outside the grid: no drainage that way
        if (x >= 0 && x < nx && y >= 0 && y < ny) {
            w = W[(int64_t)y * pitch + x];
            const float zz = z[(int64_t)y * pitch + x];
            if (!(isfinite(zz) && zz > nodata)) w = INFINITY;     // nodata cells are walls (never popped, :372-378)
        }
        sw[ly][lx] = w;
    }
    for (int k = tid; k < FT * FT; k += FT * 8) {
        const int ly = k / FT, lx = k % FT;
        const int x = x0 + lx, y = y0 + ly;
        sz[ly][lx] = (x < nx && y < ny) ? z[(int64_t)y * pitch + x] : INFINITY;
    }
    if (tid == 0) s_any = 0;
    __syncthreads();
    for (int it = 0; it < FILL_LOCAL_ITERS; ++it) {
        if (tid == 0) s_changed = 0;
        __syncthreads();
        int mine = 0;
        for (int r = threadIdx.y; r < FT; r += 8) {
            const int lx = threadIdx.x + 1, ly = r + 1;
            const float zc = sz[r][threadIdx.x];
            const float cur = sw[ly][lx];
            if (isfinite(zc) && zc > nodata && cur > zc) {
                float m = fminf(fminf(fminf(sw[ly - 1][lx - 1], sw[ly - 1][lx]), fminf(sw[ly - 1][lx + 1], sw[ly][lx - 1])),
                                fminf(fminf(sw[ly][lx + 1], sw[ly + 1][lx - 1]), fminf(sw[ly + 1][lx], sw[ly + 1][lx + 1])));
                const float nw = fmaxf(zc, m);
                if (nw < cur) { sw[ly][lx] = nw; mine = 1; }          // monotone: races only delay convergence
            }
        }
        if (mine) s_changed = 1;
        __syncthreads();
        if (!s_changed) break;
        if (tid == 0) s_any = 1;
        __syncthreads();
    }
    __syncthreads();
    if (s_any) {
        for (int k = tid; k < FT * FT; k += FT * 8) {
            const int ly = k / FT, lx = k % FT;
            const int x = x0 + lx, y = y0 + ly;
            if (x < nx && y < ny) {
                const float zc = sz[ly][lx];
                if (isfinite(zc) && zc > nodata) W[(int64_t)y * pitch + x] = sw[ly + 1][lx + 1];
            }
        }
        if (tid == 0) *changed = 1;
    }
}

__global__ void __launch_bounds__(256)
k_fill_count(int nx, int ny, int pitch, const float* __restrict__ z, const float* __restrict__ W,
             unsigned long long* counters) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    unsigned raised = 0;
    if (x < nx && y < ny) raised = (W[(int64_t)y * pitch + x] > z[(int64_t)y * pitch + x]);
    const unsigned b = __ballot_sync(0xffffffffu, raised);
    if (((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0 && b)
        atomicAdd(&counters[9], (unsigned long long)__popc(b));
}

extern "C" int tf_fill_pits(tf_handle h, const float* dem, float nodata, float* filled_out,
                            int64_t* n_raised_out, int64_t* n_sweeps_out) {
    TF_TRY(use_device(h));
    TF_CHECK(dem && filled_out, "tf_fill_pits: NULL grid");
    const Geom& g = h->g;
    TF_CHECK(!g.has_north && !g.has_south, "tf_fill_pits: the grid must be one strip (a fill is global)");
    Raster<float> Z, W;
    TF_TRY(raster_alloc(h, Z, 0));
    int rc = raster_alloc(h, W, 0);
    int* d_changed = reinterpret_cast<int*>(h->counters + 10);
    long long sweeps = 0;
    do {
        if (rc) break;
        if ((rc = raster_upload(h, Z, dem))) break;
        dim3 blk(64, 4), grd((g.nx + 63) / 64, (g.ny + 3) / 4);
        k_fill_init<<<grd, blk, 0, h->stream>>>(g.nx, g.ny, g.pitch, Z.p, nodata, W.p);
        dim3 tb(FT, 8), tg((g.nx + FT - 1) / FT, (g.ny + FT - 1) / FT);
        while (true) {
            int changed = 0;
            // a batch of passes between two looks at the flag: the flag only says "not yet done"
            if (cudaMemsetAsync(d_changed, 0, sizeof(int), h->stream) != cudaSuccess) { rc = 1; break; }
            for (int k = 0; k < 8; ++k)
                k_fill_relax<<<tg, tb, 0, h->stream>>>(g.nx, g.ny, g.pitch, Z.p, nodata, W.p, d_changed);
            sweeps += 8;
            if (cudaMemcpyAsync(&changed, d_changed, sizeof(int), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
                cudaStreamSynchronize(h->stream) != cudaSuccess) {
                tf_set_error("tf_fill_pits: %s", cudaGetErrorString(cudaGetLastError()));
                rc = 1;
                break;
            }
            if (!changed) break;
        }
        if (rc) break;
        unsigned long long n_raised = 0;
        cudaMemsetAsync(h->counters + 9, 0, sizeof(unsigned long long), h->stream);
        k_fill_count<<<grd, blk, 0, h->stream>>>(g.nx, g.ny, g.pitch, Z.p, W.p, h->counters);
        cudaMemcpyAsync(&n_raised, h->counters + 9, sizeof(n_raised), cudaMemcpyDeviceToHost, h->stream);
        if ((rc = raster_download<float>(h, W.p, filled_out))) break;
        if (n_raised_out) *n_raised_out = (int64_t)n_raised;
        if (n_sweeps_out) *n_sweeps_out = sweeps;
    } while (0);
    raster_free(Z);
    raster_free(W);
    return rc;
}

// ---------------------------------------------------------------------------
// saturated zone: water-table change from recharge and lateral D8 flow
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_satzone_dzw(int nx, int ny, int pitch, const uint8_t* __restrict__ inmask, const double* __restrict__ Q_gw,
              const double* __restrict__ Rg_grid, double Rg_scalar, const double* __restrict__ da_rows,
              int da_uniform, double dt, double* __restrict__ dzw) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= nx || y >= ny) return;
    const int64_t i = (int64_t)y * pitch + x;
    const double da = da_rows[y];
    const double Rg = Rg_grid ? Rg_grid[i] : Rg_scalar;
    double acc = dt * (Rg - (Q_gw[i] / da));                               // satzone_base.py:934
    const double factor = dt / da;                                         // :977
    const int m = inmask[i];
    // bit j: the neighbour opposite to direction j drains into this cell = the
    // k = j + 1 term of :978-1008, added in that order
    const int dr[8] = {-1, 0, 1, 1, 1, 0, -1, -1}, dc[8] = {1, 1, 1, 0, -1, -1, -1, 0};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        if ((m >> j) & 1) {
            const double q = Q_gw[i - ((int64_t)dr[j] * pitch + dc[j])];
            acc += da_uniform ? (q * factor) : ((dt * q) / da);            // :979-992 / :994-1008
        }
    }
    dzw[i] = acc;
}

extern "C" int tf_satzone_dzw(tf_handle h, const double* Q_gw, const double* Rg_grid, double Rg_scalar,
                              const double* da_rows, double dt, double* dzw_out) {
    TF_TRY(use_device(h));
    TF_CHECK(h->have_d8, "tf_satzone_dzw: build the D8 grid first");
    TF_CHECK(Q_gw && da_rows && dzw_out, "tf_satzone_dzw: NULL input");
    const Geom& g = h->g;
    Raster<double> Q, R, Z;
    double* d_da = nullptr;
    int rc = 0;
    do {
        if ((rc = raster_alloc(h, Q, 0))) break;
        if ((rc = raster_alloc(h, Z, 0))) break;
        if ((rc = raster_upload(h, Q, Q_gw))) break;
        if (Rg_grid) { if ((rc = raster_alloc(h, R, 0))) break; if ((rc = raster_upload(h, R, Rg_grid))) break; }
        if (cudaMalloc((void**)&d_da, sizeof(double) * g.ny) != cudaSuccess) { tf_set_error("tf_satzone_dzw: cudaMalloc"); rc = 1; break; }
        cudaMemcpyAsync(d_da, da_rows, sizeof(double) * g.ny, cudaMemcpyHostToDevice, h->stream);
        int uniform = 1;
        for (int y = 1; y < g.ny; ++y) uniform = uniform && (da_rows[y] == da_rows[0]);
        dim3 blk(64, 4), grd((g.nx + 63) / 64, (g.ny + 3) / 4);
        k_satzone_dzw<<<grd, blk, 0, h->stream>>>(g.nx, g.ny, g.pitch, h->inmask.p, Q.p, Rg_grid ? R.p : nullptr,
                                                  Rg_scalar, d_da, uniform, dt, Z.p);
        if (cudaGetLastError() != cudaSuccess) { tf_set_error("k_satzone_dzw launch failed"); rc = 1; break; }
        rc = raster_download<double>(h, Z.p, dzw_out);
    } while (0);
    if (d_da) cudaFree(d_da);
    raster_free(Q); raster_free(R); raster_free(Z);
    return rc;
}
