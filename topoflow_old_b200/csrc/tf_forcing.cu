// tf_forcing.cu -- the excess-rain providers on the device.
//
// In the reference the six fluxes of update_R come from other components
// (meteorology, snow, evaporation, infiltration), each a set of whole-grid
// NumPy expressions run at its own time step, and EMELI carries their outputs
// into the channels component through host arrays every channel step.  Here
// the same expressions run as kernels on device-resident grids, so that a
// coupled run never moves a grid across PCIe per step:
//
//   tf_grid_*                    device grids a caller can create, fill, read
//   tf_prov_met_rain             P_rain = P * (T_air > 0)       met_base.py:1013-1038
//   tf_prov_snow_degree_day      SM, h_swe, h_snow               snow_degree_day.py:272-312,
//                                                                snow_base.py:176-233, 365-440
//   tf_prov_evap_priestley_taylor  ET                            evap_priestley_taylor.py:290-353
//   tf_prov_infil_green_ampt     IN, I                           infil_green_ampt.py:516-572,
//                                                                infil_base.py:216-264, 500-513,
//                                                                592-647, 700-728, 784-823
//   tf_interp_linear             EMELI's time interpolation      time_interpolation.py:91-133
//   tf_grid_axpb                 a * t + b                       time_interpolation.py:677
//   tf_chan_set_forcing_dev      hand a device grid (and its rate of change)
//                                to update_R                     channels_base.py:1103-1147
//
// Every operand is a device grid (id >= 0) or a scalar (id < 0), as every
// variable of the reference is "Scalar" or "Grid" typed.  Expressions keep
// NumPy's order of operations; the TU is built with -fmad=false.
#include <math.h>

#include "tf_chan_host.cuh"

struct Opd {            // one operand of a provider expression
    const double* p;    // cell (0,0) of a device grid, or NULL
    double s;           // the scalar used when p is NULL
};
__device__ __forceinline__ double opd_at(const Opd& o, int64_t i) { return o.p ? o.p[i] : o.s; }

// np.maximum / np.minimum: NaN propagates
__device__ __forceinline__ double np_maximum(double a, double b) { return (a != a) ? a : ((b != b) ? b : (a > b ? a : b)); }
__device__ __forceinline__ double np_minimum(double a, double b) { return (a != a) ? a : ((b != b) ? b : (a < b ? a : b)); }

static int grid_ptr(tf_handle h, int gid, double** out, const char* what) {
    TF_CHECK(gid >= 0 && gid < (int)h->grids.size() && h->grids[gid].p, "%s: no device grid %d", what, gid);
    *out = h->grids[gid].p;
    return 0;
}
static int make_opd(tf_handle h, int gid, double scalar, Opd* o, const char* what) {
    o->p = nullptr; o->s = scalar;
    if (gid >= 0) { double* p = nullptr; TF_TRY(grid_ptr(h, gid, &p, what)); o->p = p; }
    return 0;
}
static dim3 cell_block() { return dim3(64, 4); }
static dim3 cell_grid(const Geom& g) { return dim3((g.nx + 63) / 64, (g.ny + 3) / 4); }

#define CELL_INDEX()                                             \
    const int x = blockIdx.x * blockDim.x + threadIdx.x;         \
    const int y = blockIdx.y * blockDim.y + threadIdx.y;         \
    if (x >= nx || y >= ny) return;                              \
    const int64_t i = (int64_t)y * pitch + x

// ---------------------------------------------------------------------------
// device grids
// ---------------------------------------------------------------------------
extern "C" int tf_grid_new(tf_handle h, int* gid_out) {
    TF_TRY(use_device(h));
    TF_CHECK(gid_out != nullptr, "tf_grid_new: gid_out is NULL");
    Raster<double> r;
    TF_TRY(raster_alloc(h, r, 0));
    h->grids.push_back(r);
    *gid_out = (int)h->grids.size() - 1;
    return 0;
}

static __global__ void k_grid_fill(int nx, int ny, int pitch, double* g, double v) {
    CELL_INDEX();
    g[i] = v;
}

extern "C" int tf_grid_set(tf_handle h, int gid, const double* host, double scalar) {
    TF_TRY(use_device(h));
    double* p = nullptr;
    TF_TRY(grid_ptr(h, gid, &p, "tf_grid_set"));
    if (host) return raster_upload(h, h->grids[gid], host);
    const Geom& g = h->g;
    k_grid_fill<<<cell_grid(g), cell_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, p, scalar);
    TF_LAUNCHED(h, h->stream, "k_grid_fill");
    return 0;
}

extern "C" int tf_grid_get(tf_handle h, int gid, double* host) {
    TF_TRY(use_device(h));
    TF_CHECK(host != nullptr, "tf_grid_get: host is NULL");
    double* p = nullptr;
    TF_TRY(grid_ptr(h, gid, &p, "tf_grid_get"));
    return raster_download<double>(h, p, host);
}

int forcing_free(tf_handle h) {
    for (auto& r : h->grids) raster_free(r);
    h->grids.clear();
    return 0;
}

// sum over the strip of (X * da) * dt, rows added north to south (the
// vol_SM / vol_ET / vol_IN increments; np.sum's pairwise order differs in the
// last bits only)
static __global__ void __launch_bounds__(256)
k_sum_flux(int nx, int pitch, const double* __restrict__ X, const double* __restrict__ da_rows, double dt,
           double* partial) {
    __shared__ double sh[256];
    const int y = blockIdx.x;
    const double da = da_rows[y];
    double acc = 0.0;
    for (int x = threadIdx.x; x < nx; x += blockDim.x) acc += (X[(int64_t)y * pitch + x] * da) * dt;
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[y] = sh[0];
}

static int ensure_partial(tf_handle h, int n) {
    if (h->partial_cap >= n) return 0;
    if (h->partial) cudaFree(h->partial);
    h->partial = nullptr;
    h->partial_cap = n;
    TF_CUDA(cudaMalloc((void**)&h->partial, sizeof(double) * n));
    return 0;
}

static int flux_volume(tf_handle h, const double* X, double dt, double* vol_out) {
    if (!vol_out) return 0;
    const Geom& g = h->g;
    TF_CHECK(h->da_rows != nullptr, "provider volume: cell areas are known after tf_chan_init only");
    TF_TRY(ensure_partial(h, 6 * g.ny));
    k_sum_flux<<<g.ny, 256, 0, h->stream>>>(g.nx, g.pitch, X, h->da_rows + TF_APRON, dt, h->partial);
    TF_LAUNCHED(h, h->stream, "k_sum_flux");
    std::vector<double> part(g.ny);
    TF_CUDA(cudaMemcpyAsync(part.data(), h->partial, sizeof(double) * g.ny, cudaMemcpyDeviceToHost, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    double s = 0.0;
    for (int y = 0; y < g.ny; ++y) s += part[y];
    *vol_out = s;
    return 0;
}

// ---------------------------------------------------------------------------
// meteorology: liquid / frozen part of the precipitation
// ---------------------------------------------------------------------------
static __global__ void k_met_rain(int nx, int ny, int pitch, Opd P, Opd T, double* P_rain, double* P_snow) {
    CELL_INDEX();
    const double p = opd_at(P, i), t = opd_at(T, i);
    if (P_rain) P_rain[i] = p * ((t > 0.0) ? 1.0 : 0.0);           // met_base.py:1037
    if (P_snow) P_snow[i] = p * ((t <= 0.0) ? 1.0 : 0.0);          // met_base.py:1081
}

extern "C" int tf_prov_met_rain(tf_handle h, int P_gid, double P, int T_air_gid, double T_air,
                                int P_rain_gid, int P_snow_gid) {
    TF_TRY(use_device(h));
    Opd oP, oT;
    TF_TRY(make_opd(h, P_gid, P, &oP, "tf_prov_met_rain"));
    TF_TRY(make_opd(h, T_air_gid, T_air, &oT, "tf_prov_met_rain"));
    double *pr = nullptr, *ps = nullptr;
    if (P_rain_gid >= 0) TF_TRY(grid_ptr(h, P_rain_gid, &pr, "tf_prov_met_rain"));
    if (P_snow_gid >= 0) TF_TRY(grid_ptr(h, P_snow_gid, &ps, "tf_prov_met_rain"));
    const Geom& g = h->g;
    k_met_rain<<<cell_grid(g), cell_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, oP, oT, pr, ps);
    TF_LAUNCHED(h, h->stream, "k_met_rain");
    return 0;
}

// ---------------------------------------------------------------------------
// snow, degree-day method
// ---------------------------------------------------------------------------
static __global__ void k_snow_degree_day(int nx, int ny, int pitch, Opd T_air, Opd P_snow, Opd c0, Opd T0,
                                         double rho_H2O, Opd rho_snow, double dt,
                                         double* h_snow, double* h_swe, double* SM) {
    CELL_INDEX();
    const double M = (opd_at(c0, i) / 8.64E7) * (opd_at(T_air, i) - opd_at(T0, i));   // snow_degree_day.py:308
    double sm = np_maximum(M, 0.0);                                                    // :309
    const double ratio = rho_H2O / opd_at(rho_snow, i);                                // snow_base.py:372
    const double sm_max = (ratio / dt) * h_snow[i];                                    // :373
    sm = np_minimum(sm, sm_max);                                                       // :374
    sm = np_maximum(sm, 0.0);                                                          // :379
    SM[i] = sm;
    double swe = h_swe[i];
    swe += opd_at(P_snow, i) * dt;                                                     // :419-422
    swe -= sm * dt;                                                                    // :425-426
    swe = np_maximum(swe, 0.0);                                                        // :428
    h_swe[i] = swe;
    h_snow[i] = swe * ratio;                                                           // :432-440
}

extern "C" int tf_prov_snow_degree_day(tf_handle h, int T_air_gid, double T_air, int P_snow_gid, double P_snow,
                                       int c0_gid, double c0, int T0_gid, double T0, double rho_H2O,
                                       int rho_snow_gid, double rho_snow, double dt,
                                       int h_snow_gid, int h_swe_gid, int SM_gid, double* vol_SM_out) {
    TF_TRY(use_device(h));
    Opd oT, oP, oc, o0, or_;
    TF_TRY(make_opd(h, T_air_gid, T_air, &oT, "tf_prov_snow_degree_day"));
    TF_TRY(make_opd(h, P_snow_gid, P_snow, &oP, "tf_prov_snow_degree_day"));
    TF_TRY(make_opd(h, c0_gid, c0, &oc, "tf_prov_snow_degree_day"));
    TF_TRY(make_opd(h, T0_gid, T0, &o0, "tf_prov_snow_degree_day"));
    TF_TRY(make_opd(h, rho_snow_gid, rho_snow, &or_, "tf_prov_snow_degree_day"));
    double *hs = nullptr, *hw = nullptr, *sm = nullptr;
    TF_TRY(grid_ptr(h, h_snow_gid, &hs, "tf_prov_snow_degree_day"));
    TF_TRY(grid_ptr(h, h_swe_gid, &hw, "tf_prov_snow_degree_day"));
    TF_TRY(grid_ptr(h, SM_gid, &sm, "tf_prov_snow_degree_day"));
    TF_CHECK(dt > 0.0, "tf_prov_snow_degree_day: dt must be positive");
    const Geom& g = h->g;
    k_snow_degree_day<<<cell_grid(g), cell_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, oT, oP, oc, o0, rho_H2O,
                                                                    or_, dt, hs, hw, sm);
    TF_LAUNCHED(h, h->stream, "k_snow_degree_day");
    return flux_volume(h, sm, dt, vol_SM_out);                                         // update_SM_integral :384-393
}

// ---------------------------------------------------------------------------
// evaporation, Priestley-Taylor
// ---------------------------------------------------------------------------
static __global__ void k_evap_pt(int nx, int ny, int pitch, Opd Qn_SW, Opd Qn_LW, Opd T_air, Opd T_surf,
                                 Opd alpha, Opd K_soil, Opd soil_x, Opd T_soil_x, double* ET) {
    CELL_INDEX();
    const double Q_net = opd_at(Qn_SW, i) + opd_at(Qn_LW, i);                          // :322
    const double Qc = (opd_at(K_soil, i) * (opd_at(T_soil_x, i) - opd_at(T_surf, i))) / opd_at(soil_x, i);   // :334
    const double Qet = (opd_at(alpha, i) * (0.406 + (0.011 * opd_at(T_air, i)))) * (Q_net - Qc);            // :335
    ET[i] = np_maximum(Qet / 2.5E+9, 0.0);                                             // :349-351
}

extern "C" int tf_prov_evap_priestley_taylor(tf_handle h, int Qn_SW_gid, double Qn_SW, int Qn_LW_gid, double Qn_LW,
                                             int T_air_gid, double T_air, int T_surf_gid, double T_surf,
                                             int alpha_gid, double alpha, int K_soil_gid, double K_soil,
                                             int soil_x_gid, double soil_x, int T_soil_x_gid, double T_soil_x,
                                             double dt, int ET_gid, double* vol_ET_out) {
    TF_TRY(use_device(h));
    Opd o[8];
    const int gids[8] = {Qn_SW_gid, Qn_LW_gid, T_air_gid, T_surf_gid, alpha_gid, K_soil_gid, soil_x_gid, T_soil_x_gid};
    const double sc[8] = {Qn_SW, Qn_LW, T_air, T_surf, alpha, K_soil, soil_x, T_soil_x};
    for (int k = 0; k < 8; ++k) TF_TRY(make_opd(h, gids[k], sc[k], &o[k], "tf_prov_evap_priestley_taylor"));
    double* et = nullptr;
    TF_TRY(grid_ptr(h, ET_gid, &et, "tf_prov_evap_priestley_taylor"));
    const Geom& g = h->g;
    k_evap_pt<<<cell_grid(g), cell_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, o[0], o[1], o[2], o[3], o[4], o[5],
                                                            o[6], o[7], et);
    TF_LAUNCHED(h, h->stream, "k_evap_pt");
    return flux_volume(h, et, dt, vol_ET_out);                                         // evap_base.py:300-309
}

// ---------------------------------------------------------------------------
// infiltration, Green-Ampt (one layer)
// ---------------------------------------------------------------------------
static __global__ void k_infil_green_ampt(int nx, int ny, int pitch, Opd P_rain, Opd SM, Opd Ks, Opd Ki, Opd qs,
                                          Opd qi, Opd G, const double* h_table, const double* elev, double dt,
                                          double* I, double* IN) {
    CELL_INDEX();
    const double P_total = opd_at(P_rain, i) + opd_at(SM, i);                          // infil_base.py:509
    const double ks = opd_at(Ks, i);
    const double dK = ks - opd_at(Ki, i);                                              // infil_green_ampt.py:544
    const double dq = opd_at(qs, i) - opd_at(qi, i);                                   // :545
    const double fc = (((opd_at(G, i) * dK) * dq) / I[i]) + ks;                        // :546
    double in = np_minimum(fc, P_total);                                               // :567
    if (P_total < ks) in = P_total;                                                    // check_low_rainrate :805-822
    if (h_table && elev && h_table[i] >= elev[i]) in = 0.0;                            // adjust_infil_rate :636-639
    IN[i] = in;
    I[i] += in * dt;                                                                   // update_I :711
}

extern "C" int tf_prov_infil_green_ampt(tf_handle h, int P_rain_gid, double P_rain, int SM_gid, double SM,
                                        int Ks_gid, double Ks, int Ki_gid, double Ki, int qs_gid, double qs,
                                        int qi_gid, double qi, int G_gid, double G, int h_table_gid, int elev_gid,
                                        double dt, int I_gid, int IN_gid, double* vol_IN_out) {
    TF_TRY(use_device(h));
    Opd o[7];
    const int gids[7] = {P_rain_gid, SM_gid, Ks_gid, Ki_gid, qs_gid, qi_gid, G_gid};
    const double sc[7] = {P_rain, SM, Ks, Ki, qs, qi, G};
    for (int k = 0; k < 7; ++k) TF_TRY(make_opd(h, gids[k], sc[k], &o[k], "tf_prov_infil_green_ampt"));
    double *I = nullptr, *IN = nullptr, *ht = nullptr, *el = nullptr;
    TF_TRY(grid_ptr(h, I_gid, &I, "tf_prov_infil_green_ampt"));
    TF_TRY(grid_ptr(h, IN_gid, &IN, "tf_prov_infil_green_ampt"));
    if (h_table_gid >= 0 && elev_gid >= 0) {
        TF_TRY(grid_ptr(h, h_table_gid, &ht, "tf_prov_infil_green_ampt"));
        TF_TRY(grid_ptr(h, elev_gid, &el, "tf_prov_infil_green_ampt"));
    }
    TF_CHECK(dt > 0.0, "tf_prov_infil_green_ampt: dt must be positive");
    const Geom& g = h->g;
    k_infil_green_ampt<<<cell_grid(g), cell_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, o[0], o[1], o[2], o[3],
                                                                     o[4], o[5], o[6], ht, el, dt, I, IN);
    TF_LAUNCHED(h, h->stream, "k_infil_green_ampt");
    return flux_volume(h, IN, dt, vol_IN_out);                                         // update_IN_integral :650-660
}

// ---------------------------------------------------------------------------
// EMELI's linear time interpolation of a provided grid
// (time_interpolation.py:91-133): after a provider's update,
//     dv = |v2 - v1|;  if dv.min() != 0 and t2 != t1:
//         a = (v2 - v1) / (t2 - t1);  b = v2 - a * t2      else a = 0, b = v2
// -- interpolation only when EVERY cell changed, a step function otherwise.
// ---------------------------------------------------------------------------
static __global__ void __launch_bounds__(256)
k_absdiff_min(int nx, int pitch, const double* __restrict__ v1, const double* __restrict__ v2, double* partial) {
    // partial[y] = min over the row of |v2 - v1| (NaN if any NaN, as ndarray.min)
    __shared__ double sh[256];
    const int y = blockIdx.x;
    double m = INFINITY;
    bool nan_seen = false;
    for (int x = threadIdx.x; x < nx; x += blockDim.x) {
        const double d = fabs(v2[(int64_t)y * pitch + x] - v1[(int64_t)y * pitch + x]);
        if (d != d) nan_seen = true;
        m = fmin(m, d);
    }
    sh[threadIdx.x] = nan_seen ? NAN : m;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            const double p = sh[threadIdx.x], q = sh[threadIdx.x + s];
            sh[threadIdx.x] = (p != p || q != q) ? NAN : (p < q ? p : q);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[y] = sh[0];
}

static __global__ void k_interp_ab(int nx, int ny, int pitch, const double* __restrict__ v1,
                                   const double* __restrict__ v2, double t2, double dt21, int linear,
                                   double* a, double* b) {
    CELL_INDEX();
    const double x2 = v2[i];
    if (linear) {
        const double aa = (x2 - v1[i]) / dt21;
        a[i] = aa;
        b[i] = x2 - (aa * t2);
    } else {
        a[i] = 0.0;
        b[i] = x2;
    }
}

extern "C" int tf_interp_linear(tf_handle h, int v1_gid, int v2_gid, double t1, double t2, int a_gid, int b_gid,
                                int* linear_out) {
    TF_TRY(use_device(h));
    double *v1 = nullptr, *v2 = nullptr, *a = nullptr, *b = nullptr;
    TF_TRY(grid_ptr(h, v1_gid, &v1, "tf_interp_linear"));
    TF_TRY(grid_ptr(h, v2_gid, &v2, "tf_interp_linear"));
    TF_TRY(grid_ptr(h, a_gid, &a, "tf_interp_linear"));
    TF_TRY(grid_ptr(h, b_gid, &b, "tf_interp_linear"));
    const Geom& g = h->g;
    TF_TRY(ensure_partial(h, 6 * g.ny));
    k_absdiff_min<<<g.ny, 256, 0, h->stream>>>(g.nx, g.pitch, v1, v2, h->partial);
    TF_LAUNCHED(h, h->stream, "k_absdiff_min");
    std::vector<double> part(g.ny);
    TF_CUDA(cudaMemcpyAsync(part.data(), h->partial, sizeof(double) * g.ny, cudaMemcpyDeviceToHost, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    double m = INFINITY;
    bool nan_seen = false;
    for (int y = 0; y < g.ny; ++y) { if (part[y] != part[y]) nan_seen = true; else if (part[y] < m) m = part[y]; }
    // several strips: the minimum is the grid's (one 8-byte all-reduce of a flag per strip)
    unsigned long long zero_here = (!nan_seen && m == 0.0) ? 1ull : 0ull;
    if (h->nccl_comm) {
        TF_CUDA(cudaMemcpyAsync(h->counters + 12, &zero_here, 8, cudaMemcpyHostToDevice, h->stream));
        TF_TRY(comm_allreduce_u64(h, h->stream, h->counters + 12, 1));
        TF_CUDA(cudaMemcpyAsync(&zero_here, h->counters + 12, 8, cudaMemcpyDeviceToHost, h->stream));
        TF_CUDA(cudaStreamSynchronize(h->stream));
    }
    // "dv_min != 0" is also true for NaN, as in NumPy
    const int linear = (zero_here == 0ull) && (t2 != t1);
    k_interp_ab<<<cell_grid(g), cell_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, v1, v2, t2, t2 - t1, linear, a, b);
    TF_LAUNCHED(h, h->stream, "k_interp_ab");
    if (linear_out) *linear_out = linear;
    return 0;
}

static __global__ void k_axpb(int nx, int ny, int pitch, Opd a, Opd b, double t, double* out) {
    CELL_INDEX();
    out[i] = (opd_at(a, i) * t) + opd_at(b, i);                    // time_interpolation.py:677
}

extern "C" int tf_grid_axpb(tf_handle h, int a_gid, double a, int b_gid, double b, double t, int out_gid) {
    TF_TRY(use_device(h));
    Opd oa, ob;
    TF_TRY(make_opd(h, a_gid, a, &oa, "tf_grid_axpb"));
    TF_TRY(make_opd(h, b_gid, b, &ob, "tf_grid_axpb"));
    double* out = nullptr;
    TF_TRY(grid_ptr(h, out_gid, &out, "tf_grid_axpb"));
    const Geom& g = h->g;
    k_axpb<<<cell_grid(g), cell_block(), 0, h->stream>>>(g.nx, g.ny, g.pitch, oa, ob, t, out);
    TF_LAUNCHED(h, h->stream, "k_axpb");
    return 0;
}

extern "C" int tf_grid_copy(tf_handle h, int src_gid, int dst_gid) {
    TF_TRY(use_device(h));
    double *s = nullptr, *d = nullptr;
    TF_TRY(grid_ptr(h, src_gid, &s, "tf_grid_copy"));
    TF_TRY(grid_ptr(h, dst_gid, &d, "tf_grid_copy"));
    TF_CUDA(cudaMemcpyAsync(h->grids[dst_gid].base, h->grids[src_gid].base, sizeof(double) * (size_t)h->g.n_alloc,
                            cudaMemcpyDeviceToDevice, h->stream));
    return 0;
}

// ---------------------------------------------------------------------------
// update_R from device-resident fluxes: flux k of R = (P_rain + SM + GW + MR)
// - (ET + IN) is value_k + rate_k * t, each a device grid or a scalar.
// tf_chan_assemble_R_dev forms R0 = the combination of the values and, when a
// rate is non-zero, R1 = the same combination of the rates; the step kernels
// then use R = R1 * t + R0 with t the channel time of the step -- no grid is
// uploaded, none is re-assembled between provider updates.
// ---------------------------------------------------------------------------
extern "C" int tf_chan_set_forcing_dev(tf_handle h, int which, int value_gid, double value, int rate_gid,
                                       double rate) {
    TF_TRY(use_device(h));
    TF_CHECK(h->chan_ready, "tf_chan_set_forcing_dev: call tf_chan_init first");
    TF_CHECK(which >= 0 && which < TF_R_COUNT, "tf_chan_set_forcing_dev: unknown flux %d", which);
    Opd v, r;
    TF_TRY(make_opd(h, value_gid, value, &v, "tf_chan_set_forcing_dev"));
    TF_TRY(make_opd(h, rate_gid, rate, &r, "tf_chan_set_forcing_dev"));
    h->fdev_value[which] = v.p; h->fdev_value_s[which] = v.s;
    h->fdev_rate[which] = r.p; h->fdev_rate_s[which] = r.s;
    return 0;
}

static int row_partials_sum(tf_handle h, double* out) {
    const Geom& g = h->g;
    std::vector<double> part(g.ny);
    TF_CUDA(cudaMemcpyAsync(part.data(), h->partial, sizeof(double) * g.ny, cudaMemcpyDeviceToHost, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    double s = 0.0;
    for (int y = 0; y < g.ny; ++y) s += part[y];
    *out = s;
    return 0;
}

// sum0_out / sum1_out (optional): sum over the strip of (R0*da)*dt and of
// (R1*da)*dt, so that the caller can keep update_R_integral
// (channels_base.py:1151-1160) per step as sum1 * t + sum0.
extern "C" int tf_chan_assemble_R_dev(tf_handle h, double* sum0_out, double* sum1_out) {
    TF_TRY(use_device(h));
    TF_CHECK(h->chan_ready, "tf_chan_assemble_R_dev: call tf_chan_init first");
    const Geom& g = h->g;
    TF_TRY(raster_alloc(h, h->R, 0));
    TF_TRY(ensure_partial(h, 6 * g.ny));
    ChanArgs a = chan_args(h, 0, g.ny);
    ForcingArgs fa;
    bool any_rate = false;
    for (int k = 0; k < TF_R_COUNT; ++k) {
        fa.grid[k] = h->fdev_value[k]; fa.scalar[k] = h->fdev_value_s[k];
        any_rate = any_rate || h->fdev_rate[k] || h->fdev_rate_s[k] != 0.0;
    }
    k_assemble_R<<<g.ny, 256, 0, h->stream>>>(a, fa, h->R.p, h->partial);
    TF_LAUNCHED(h, h->stream, "k_assemble_R");
    h->R_is_grid = true;
    h->R_rate_on = false;
    if (sum0_out) TF_TRY(row_partials_sum(h, sum0_out));
    if (sum1_out) *sum1_out = 0.0;
    if (any_rate) {
        TF_TRY(raster_alloc(h, h->R1, 0));
        for (int k = 0; k < TF_R_COUNT; ++k) { fa.grid[k] = h->fdev_rate[k]; fa.scalar[k] = h->fdev_rate_s[k]; }
        k_assemble_R<<<g.ny, 256, 0, h->stream>>>(a, fa, h->R1.p, h->partial);
        TF_LAUNCHED(h, h->stream, "k_assemble_R");
        h->R_rate_on = true;
        if (sum1_out) TF_TRY(row_partials_sum(h, sum1_out));
    }
    return 0;
}
