// tf_chan.cu -- host side of the channel-routing path: device layout, the
// per-step launch sequence and the C ABI around it.
#include <stdlib.h>

#include "tf_chan_host.cuh"

int chan_layout_init(tf_handle h, int allow_packed);   // tf_chan_layout.cu
int chan_exchange_static(tf_handle h);                  // tf_chan_layout.cu

// strips + marching step: phase A of the apron row reads the neighbour's R
static int chan_exchange_R(tf_handle h) {
    if (!h->nccl_comm || !h->use_march || h->wave == TF_KINEMATIC) return 0;
    void* arr[1] = {h->R.p};
    size_t es[1] = {sizeof(double)};
    return comm_exchange_rows(h, h->stream, arr, es, 1);
}

// ---------------------------------------------------------------------------
extern "C" int tf_chan_set_rho(tf_handle h, double rho) {
    TF_CHECK(h != nullptr, "null handle");
    h->rho = rho;
    return 0;
}

extern "C" int tf_chan_init(tf_handle h, int wave, int law, double dt, const double* da_rows,
                            const int32_t* outlet_rc, int n_outlets, int allow_packed) {
    TF_TRY(use_device(h));
    TF_CHECK(h->have_d8, "tf_chan_init: build the D8 grid first (tf_d8_build / tf_d8_set_codes)");
    TF_CHECK(!h->lr_periodic && !h->tb_periodic,
             "tf_chan_init: periodic flow grids are not routable (the reference's "
             "update_parent_IDs cannot handle them either, d8_global.py:650-655)");
    TF_CHECK(wave >= 0 && wave <= 2, "tf_chan_init: unknown wave type %d", wave);
    TF_CHECK(law == 0 || law == 1, "tf_chan_init: unknown friction law %d", law);
    TF_CHECK(dt > 0.0, "tf_chan_init: dt must be positive");
    TF_CHECK(da_rows != nullptr, "tf_chan_init: da_rows is NULL");
    for (int k = 0; k < TF_P_COUNT; ++k)
        TF_CHECK(h->hp[k].set, "tf_chan_init: parameter %d was not set", k);
    const Geom& g = h->g;
    h->wave = wave; h->law = law; h->dt = dt;
    // only grid-typed width / angle are rewritten (channels_base.py:1451-1454)
    h->mut_width = h->hp[TF_P_WIDTH].is_grid;
    h->mut_angle = h->hp[TF_P_TAN_ANGLE].is_grid;
    // cell areas per row, with apron slots
    const int n_da = g.ny + 2 * TF_APRON;
    if (!h->da_rows) TF_CUDA(cudaMalloc((void**)&h->da_rows, sizeof(double) * n_da));
    {
        // apron slots start as copies of the edge rows; a strip with neighbours
        // gets the neighbours' values in chan_exchange_static (tf_chan_layout.cu)
        std::vector<double> tmp(n_da, 0.0);
        for (int y = -TF_APRON; y < g.ny + TF_APRON; ++y)
            tmp[y + TF_APRON] = da_rows[y < 0 ? 0 : (y >= g.ny ? g.ny - 1 : y)];
        TF_CUDA(cudaMemcpy(h->da_rows, tmp.data(), sizeof(double) * n_da, cudaMemcpyHostToDevice));
    }
    // state rasters
    TF_TRY(raster_alloc(h, h->vol, 0));
    TF_TRY(raster_alloc(h, h->d, 0));
    TF_TRY(raster_alloc(h, h->u, 0));
    TF_TRY(raster_alloc(h, h->Qa, 0));
    TF_TRY(raster_alloc(h, h->Qb, 0));
    TF_TRY(raster_alloc(h, h->mstate, 0));
    if (wave == TF_DYNAMIC) TF_TRY(raster_alloc(h, h->u2, 0));
    // A handle may be initialised again (new parameters, a second run): the
    // reference zeroes u, Q and R and restarts vol / d from d0 at every
    // initialize() (channels_base.py:900-910), so nothing of the last run may
    // survive in rasters that raster_alloc found already allocated.
    {
        const size_t n64 = sizeof(double) * (size_t)g.n_alloc;
        Raster<double>* st[] = {&h->vol, &h->d, &h->u, &h->u2, &h->Qa, &h->Qb};
        for (Raster<double>* r : st)
            if (r->base) TF_CUDA(cudaMemsetAsync(r->base, 0, n64, h->stream));
        TF_CUDA(cudaMemsetAsync(h->mstate.base, 0, (size_t)g.n_alloc, h->stream));
    }
    h->R_is_grid = false;
    h->R_scalar = 0.0;
    h->R_rate_on = false;
    for (int k = 0; k < TF_R_COUNT; ++k) {
        h->forc_is_grid[k] = false; h->forc_scalar[k] = 0.0;
        h->fdev_value[k] = h->fdev_rate[k] = nullptr; h->fdev_value_s[k] = h->fdev_rate_s[k] = 0.0;
    }
    h->Qcur = h->Qa.p; h->Qnext = h->Qb.p;
    TF_TRY(chan_layout_init(h, allow_packed));
    // The marching kernels (tf_chan_march.cuh) are opt-in, TF_B200_MARCH=1: they
    // are bit-identical to the tile kernels and move fewer bytes for the
    // diffusive / dynamic wave, but as measured on B200 they are slower (DESIGN.md
    // section 3b).  They need the packed layout with ds / dw derived from the code.
    { const char* e = getenv("TF_B200_MARCH");
      h->use_march = h->packed && h->len_rows != nullptr && h->rcp_ok && (e && e[0] == '1'); }
    { const char* e = getenv("TF_B200_SEG_ROWS"); h->seg_rows = (e && atoi(e) > 0) ? atoi(e) : 64; }
    if (h->use_march && wave != TF_KINEMATIC) {
        TF_TRY(raster_alloc(h, h->volB, 0));
        TF_CUDA(cudaMemsetAsync(h->volB.base, 0, sizeof(double) * (size_t)g.n_alloc, h->stream));
    }
    if (h->use_march) TF_TRY(chan_exchange_static(h));
    h->ghost_dirty = true;
    // outlets
    TF_CHECK(n_outlets >= 0 && (n_outlets == 0 || outlet_rc), "tf_chan_init: bad outlet list");
    h->n_outlets = n_outlets;
    h->outlet_idx_h.resize(n_outlets);
    for (int k = 0; k < n_outlets; ++k) {
        const int r = outlet_rc[2 * k], c = outlet_rc[2 * k + 1];
        TF_CHECK(r >= 0 && r < g.ny && c >= 0 && c < g.nx,
                 "tf_chan_init: outlet %d (row %d, col %d) lies outside the %d x %d strip", k, r, c, g.ny, g.nx);
        h->outlet_idx_h[k] = (int64_t)r * g.pitch + c;
    }
    if (h->outlet_idx) { cudaFree(h->outlet_idx); h->outlet_idx = nullptr; }
    if (n_outlets) {
        TF_CUDA(cudaMalloc((void**)&h->outlet_idx, sizeof(int64_t) * n_outlets));
        TF_CUDA(cudaMemcpy(h->outlet_idx, h->outlet_idx_h.data(), sizeof(int64_t) * n_outlets, cudaMemcpyHostToDevice));
    }
    const long long none = 0x7fffffffffffffffLL;
    const long long init_bad[2] = {none, none};
    TF_CUDA(cudaMemcpy(h->bad, init_bad, sizeof(init_bad), cudaMemcpyHostToDevice));
    h->chan_ready = true;
    h->state_dirty = true;
    h->steps_done = 0;
    h->acc_ms = 0.0; h->acc_steps = 0; h->acc_launches = 0; h->timing_open = false;
    TF_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int tf_chan_set_state(tf_handle h, int which, const double* grid) {
    TF_TRY(use_device(h));
    TF_CHECK(h->chan_ready, "tf_chan_set_state: call tf_chan_init first");
    TF_CHECK(grid != nullptr, "tf_chan_set_state: grid is NULL");
    Raster<double> r;
    switch (which) {
        case TF_F_VOL: r = h->vol; break;
        case TF_F_D: r = h->d; h->state_dirty = true; break;
        case TF_F_U: r = h->u; break;
        case TF_F_QNEXT: r.p = h->Qnext; break;
        case TF_F_Q: r.p = h->Qcur; break;
        default:
            tf_set_error("tf_chan_set_state: field %d is not a stored state grid", which);
            return 1;
    }
    h->ghost_dirty = true;
    return raster_upload(h, r, grid);
}

extern "C" int tf_chan_set_R(tf_handle h, const double* grid, double scalar, double* sum_out) {
    TF_TRY(use_device(h));
    TF_CHECK(h->chan_ready, "tf_chan_set_R: call tf_chan_init first");
    const Geom& g = h->g;
    if (!grid) {
        h->R_is_grid = false;
        h->R_rate_on = false;
        h->R_scalar = scalar;
        if (sum_out) {
            // (R*da*dt) * n_pixels for a scalar R with scalar da (:1157-1158);
            // per-row da sums the same product row by row
            std::vector<double> da(g.ny);
            TF_CUDA(cudaMemcpy(da.data(), h->da_rows + TF_APRON, sizeof(double) * g.ny, cudaMemcpyDeviceToHost));
            bool uniform = true;
            for (int y = 0; y < g.ny; ++y) uniform = uniform && (da[y] == da[0]);
            if (uniform) *sum_out = ((scalar * da[0]) * h->dt) * ((double)g.nx * g.ny);
            else {
                double s = 0.0;
                for (int y = 0; y < g.ny; ++y) s += ((scalar * da[y]) * h->dt) * g.nx;
                *sum_out = s;
            }
        }
        return 0;
    }
    TF_TRY(raster_alloc(h, h->R, 0));
    TF_TRY(raster_upload(h, h->R, grid));
    h->R_is_grid = true;
    h->R_rate_on = false;
    TF_TRY(chan_exchange_R(h));
    if (sum_out) {
        if (h->partial_cap < 6 * g.ny) {
            if (h->partial) cudaFree(h->partial);
            h->partial_cap = 6 * g.ny;
            TF_CUDA(cudaMalloc((void**)&h->partial, sizeof(double) * h->partial_cap));
        }
        ChanArgs a = chan_args(h, 0, g.ny);
        k_sum_R<<<g.ny, 256, 0, h->stream>>>(a, h->partial);
        TF_CUDA(cudaGetLastError());
        std::vector<double> part(g.ny);
        TF_CUDA(cudaMemcpyAsync(part.data(), h->partial, sizeof(double) * g.ny, cudaMemcpyDeviceToHost, h->stream));
        TF_CUDA(cudaStreamSynchronize(h->stream));
        double s = 0.0;
        for (int y = 0; y < g.ny; ++y) s += part[y];
        *sum_out = s;
    }
    return 0;
}

extern "C" int tf_chan_set_forcing(tf_handle h, int which, const double* grid, double scalar) {
    TF_TRY(use_device(h));
    TF_CHECK(h->chan_ready, "tf_chan_set_forcing: call tf_chan_init first");
    TF_CHECK(which >= 0 && which < TF_R_COUNT, "tf_chan_set_forcing: unknown flux %d", which);
    if (!grid) {
        h->forc_is_grid[which] = false;
        h->forc_scalar[which] = scalar;
        return 0;
    }
    TF_TRY(raster_alloc(h, h->forc[which], 0));
    TF_TRY(raster_upload(h, h->forc[which], grid));
    h->forc_is_grid[which] = true;
    return 0;
}

extern "C" int tf_chan_assemble_R(tf_handle h, double* sum_out) {
    TF_TRY(use_device(h));
    TF_CHECK(h->chan_ready, "tf_chan_assemble_R: call tf_chan_init first");
    const Geom& g = h->g;
    bool any_grid = false;
    for (int k = 0; k < TF_R_COUNT; ++k) any_grid = any_grid || h->forc_is_grid[k];
    const double* v = h->forc_scalar;
    if (!any_grid)
        return tf_chan_set_R(h, nullptr,
                             (((v[TF_R_P_RAIN] + v[TF_R_SM]) + v[TF_R_GW]) + v[TF_R_MR])
                                 - (v[TF_R_ET] + v[TF_R_IN]), sum_out);
    TF_TRY(raster_alloc(h, h->R, 0));
    if (h->partial_cap < 6 * g.ny) {
        if (h->partial) cudaFree(h->partial);
        h->partial_cap = 6 * g.ny;
        TF_CUDA(cudaMalloc((void**)&h->partial, sizeof(double) * h->partial_cap));
    }
    ForcingArgs fa;
    for (int k = 0; k < TF_R_COUNT; ++k) {
        fa.grid[k] = h->forc_is_grid[k] ? h->forc[k].p : nullptr;
        fa.scalar[k] = h->forc_scalar[k];
    }
    ChanArgs a = chan_args(h, 0, g.ny);
    k_assemble_R<<<g.ny, 256, 0, h->stream>>>(a, fa, h->R.p, h->partial);
    TF_LAUNCHED(h, h->stream, "k_assemble_R");
    h->R_is_grid = true;
    h->R_rate_on = false;
    TF_TRY(chan_exchange_R(h));
    if (sum_out) {
        std::vector<double> part(g.ny);
        TF_CUDA(cudaMemcpyAsync(part.data(), h->partial, sizeof(double) * g.ny, cudaMemcpyDeviceToHost, h->stream));
        TF_CUDA(cudaStreamSynchronize(h->stream));
        double s = 0.0;
        for (int y = 0; y < g.ny; ++y) s += part[y];
        *sum_out = s;
    }
    return 0;
}

extern "C" int tf_chan_set_points(tf_handle h, int kind, const int32_t* rows, const int32_t* cols,
                                  const double* values, int n) {
    TF_TRY(use_device(h));
    TF_CHECK(kind >= 0 && kind < 4, "tf_chan_set_points: unknown point set %d", kind);
    TF_CHECK(n >= 0 && (n == 0 || (rows && cols && values)), "tf_chan_set_points: NULL input");
    const Geom& g = h->g;
    PointSet& ps = h->pts[kind];
    if (ps.n != n) {
        if (ps.idx) cudaFree(ps.idx);
        if (ps.val) cudaFree(ps.val);
        ps.idx = nullptr; ps.val = nullptr; ps.n = 0;
        if (n) {
            TF_CUDA(cudaMalloc((void**)&ps.idx, sizeof(int64_t) * n));
            TF_CUDA(cudaMalloc((void**)&ps.val, sizeof(double) * n));
        }
        ps.n = n;
    }
    if (n) {
        std::vector<int64_t> idx(n);
        for (int k = 0; k < n; ++k) {
            TF_CHECK(rows[k] >= 0 && rows[k] < g.ny && cols[k] >= 0 && cols[k] < g.nx,
                     "tf_chan_set_points: point %d (row %d, col %d) outside the strip", k, rows[k], cols[k]);
            idx[k] = (int64_t)rows[k] * g.pitch + cols[k];
        }
        TF_CUDA(cudaMemcpy(ps.idx, idx.data(), sizeof(int64_t) * n, cudaMemcpyHostToDevice));
        TF_CUDA(cudaMemcpy(ps.val, values, sizeof(double) * n, cudaMemcpyHostToDevice));
    }
    int nmax = 0;
    for (int k = 0; k < 4; ++k) nmax = nmax > h->pts[k].n ? nmax : h->pts[k].n;
    if (nmax > h->pts_tmp_cap) {
        if (h->pts_tmp) cudaFree(h->pts_tmp);
        TF_CUDA(cudaMalloc((void**)&h->pts_tmp, sizeof(double) * 2 * nmax));
        h->pts_tmp_cap = nmax;
    }
    return 0;
}
