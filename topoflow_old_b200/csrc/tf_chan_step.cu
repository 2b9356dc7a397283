// tf_chan_step.cu -- the per-step launch sequence and the read-back entry
// points of the routing path.
#include "tf_chan_host.cuh"

// one launch covering rows [y0, y1)
static int launch_kinematic(tf_handle h, cudaStream_t s, int y0, int y1) {
    ChanArgs a = chan_args(h, y0, y1);
    if (h->kfac.p) {
        if (h->packed) k_step_kinematic<true, 0, true><<<tile_grid(h->g, y1 - y0), dim3(32, 8), 0, s>>>(a);
        else k_step_kinematic<false, 0, true><<<tile_grid(h->g, y1 - y0), dim3(32, 8), 0, s>>>(a);
    } else {
        TF_DISPATCH(h, (k_step_kinematic<P, L, false><<<tile_grid(h->g, y1 - y0), dim3(32, 8), 0, s>>>(a)));
    }
    TF_LAUNCHED(h, s, "k_step_kinematic");
    h->acc_launches++;
    return 0;
}

static int launch_phase_a(tf_handle h, cudaStream_t s, int y0, int y1) {
    ChanArgs a = chan_args(h, y0, y1);
    if (h->packed) k_phase_a<true><<<tile_grid(h->g, y1 - y0), dim3(32, 8), 0, s>>>(a);
    else k_phase_a<false><<<tile_grid(h->g, y1 - y0), dim3(32, 8), 0, s>>>(a);
    TF_LAUNCHED(h, s, "k_phase_a");
    h->acc_launches++;
    return 0;
}

static int launch_phase_b(tf_handle h, cudaStream_t s, int y0, int y1) {
    ChanArgs a = chan_args(h, y0, y1);
    if (h->wave == TF_DIFFUSIVE)
        TF_DISPATCH(h, (k_phase_b<P, L, 1><<<tile_grid(h->g, y1 - y0), dim3(32, 8), 0, s>>>(a)));
    else
        TF_DISPATCH(h, (k_phase_b<P, L, 2><<<tile_grid(h->g, y1 - y0), dim3(32, 8), 0, s>>>(a)));
    TF_LAUNCHED(h, s, "k_phase_b");
    h->acc_launches++;
    return 0;
}

// ---- which rasters cross a strip boundary in one step of the tile kernels ----
// One table for the launch sequence below and for tf_halo_plan (which
// topoflow_old_b200/strips.py and the CPU strip test read): per wave, the
// fields (enum tf_field) exchanged between the two phases and at the end of
// the step, -1 terminated.
static const int k_halo_plan[3][2][3] = {
    /* kinematic */ {{-1, -1, -1}, {TF_F_QNEXT, -1, -1}},
    /* diffusive */ {{TF_F_D, -1, -1}, {TF_F_QNEXT, -1, -1}},
    /* dynamic   */ {{TF_F_D, -1, -1}, {TF_F_QNEXT, TF_F_U, -1}}};

extern "C" int tf_halo_plan(int wave, int point, int* fields_out3) {
    TF_CHECK(wave >= 0 && wave <= 2 && (point == 0 || point == 1) && fields_out3, "tf_halo_plan: bad argument");
    for (int k = 0; k < 3; ++k) fields_out3[k] = k_halo_plan[wave][point][k];
    return 0;
}

static int exchange_planned(tf_handle h, cudaStream_t s, int point) {
    void* arr[3];
    size_t es[3];
    int n = 0;
    for (int k = 0; k < 3 && k_halo_plan[h->wave][point][k] >= 0; ++k) {
        const int f = k_halo_plan[h->wave][point][k];
        arr[n] = (f == TF_F_D) ? (void*)h->d.p : ((f == TF_F_U) ? (void*)h->u.p : (void*)h->Qnext);
        es[n++] = sizeof(double);
    }
    return n ? comm_exchange_rows(h, s, arr, es, n) : 0;
}

// ---- marching kernels (packed layout, tf_chan_march.cuh) ----------------------
template <int WAVE, int LAW, bool RG>
static int launch_march_t(tf_handle h, cudaStream_t s, int y0, int y1) {
    using C = MarchCfg<WAVE, LAW, RG>;
    static bool attr_set[64] = {false};
    if (h->device < 64 && !attr_set[h->device]) {
        TF_CUDA(cudaFuncSetAttribute(k_march<WAVE, LAW, RG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)C::SMEM));
        attr_set[h->device] = true;
    }
    ChanArgs a = chan_args(h, y0, y1);
    a.n_bands = (h->g.pitch + C::STRIDE - 1) / C::STRIDE;
    const int n_segs = (y1 - y0 + a.seg_rows - 1) / a.seg_rows;
    const long long tasks = (long long)a.n_bands * n_segs;          // one warp (= one block) each
    TF_CHECK(tasks < (1ll << 31), "tf_chan_step: %lld marching tasks; raise TF_B200_SEG_ROWS", tasks);
    k_march<WAVE, LAW, RG><<<(unsigned)tasks, 32, C::SMEM, s>>>(a);
    TF_LAUNCHED(h, s, "k_march");
    h->acc_launches++;
    return 0;
}

template <int WAVE>
static int launch_march_w(tf_handle h, cudaStream_t s, int y0, int y1) {
    if (h->law == 0) return h->R_is_grid ? launch_march_t<WAVE, 0, true>(h, s, y0, y1)
                                         : launch_march_t<WAVE, 0, false>(h, s, y0, y1);
    return h->R_is_grid ? launch_march_t<WAVE, 1, true>(h, s, y0, y1)
                        : launch_march_t<WAVE, 1, false>(h, s, y0, y1);
}

static int launch_march(tf_handle h, cudaStream_t s, int y0, int y1) {
    if (y1 <= y0) return 0;
    switch (h->wave) {
        case TF_KINEMATIC: return launch_march_w<0>(h, s, y0, y1);
        case TF_DIFFUSIVE: return launch_march_w<1>(h, s, y0, y1);
        default: return launch_march_w<2>(h, s, y0, y1);
    }
}

// one marching step: a strip with neighbours computes its edge bands first on
// the high-priority stream, so that the rows its neighbours need travel while
// the interior is computed.  One exchange per step for every wave: the next
// discharge (two rows for the fused waves, whose phase A of the apron row reads
// one row further out) and, for the dynamic wave, the new velocity.
static int march_step(tf_handle h) {
    const Geom& g = h->g;
    TF_CHECK(!h->R_rate_on, "tf_chan_step: the marching kernels do not take an interpolated R (unset TF_B200_MARCH)");
    cudaStream_t s = h->stream;
    const bool multi = h->nccl_comm != nullptr;
    const bool fused = h->wave != TF_KINEMATIC;
    void* arr[2] = {h->Qnext, h->u.p};
    size_t es[2] = {sizeof(double), sizeof(double)};
    int dep[2] = {fused ? 2 : 1, 1};
    const int n_arr = h->wave == TF_DYNAMIC ? 2 : 1;
    if (!multi || g.ny < 32) {
        TF_TRY(launch_march(h, s, 0, g.ny));
        if (multi) TF_TRY(comm_exchange_rows(h, s, arr, es, n_arr, dep));
    } else {
        cudaStream_t e = h->stream_edge;
        TF_CUDA(cudaEventRecord(h->ev_main, s));
        TF_CUDA(cudaStreamWaitEvent(e, h->ev_main, 0));
        TF_TRY(launch_march(h, e, 0, 8));
        TF_TRY(launch_march(h, e, g.ny - 8, g.ny));
        TF_TRY(comm_exchange_rows(h, e, arr, es, n_arr, dep));
        TF_CUDA(cudaEventRecord(h->ev_edge, e));
        TF_TRY(launch_march(h, s, 8, g.ny - 8));
        TF_CUDA(cudaStreamWaitEvent(s, h->ev_edge, 0));
    }
    if (fused) { Raster<double> r = h->vol; h->vol = h->volB; h->volB = r; }
    return 0;
}

// apron copies of what a neighbour strip owns and this strip's phase A of the
// apron row reads: refreshed after anything but a step changed them
static int refresh_ghosts(tf_handle h) {
    if (!h->nccl_comm) { h->ghost_dirty = false; return 0; }
    void* arr[4] = {h->vol.p, h->mstate.p, h->Qcur, h->u2.p};
    size_t es[4] = {sizeof(double), 1, sizeof(double), sizeof(double)};
    int dep[4] = {1, 1, TF_APRON, 1};
    TF_TRY(comm_exchange_rows(h, h->stream, arr, es, h->wave == TF_DYNAMIC ? 4 : 3, dep));
    h->ghost_dirty = false;
    return 0;
}

static int ensure_taps(tf_handle h, int n_steps) {
    const int nc = h->pts[TF_PT_CANALS_IN].n;
    if (n_steps <= h->taps_cap && h->taps && (nc == 0 || h->canal_taps)) return 0;
    if (h->taps) cudaFree(h->taps);
    if (h->canal_taps) cudaFree(h->canal_taps);
    h->taps = nullptr; h->canal_taps = nullptr;
    h->taps_cap = n_steps > 64 ? n_steps : 64;
    const size_t n = (size_t)h->taps_cap * (h->n_outlets > 0 ? h->n_outlets : 1) * 4;
    TF_CUDA(cudaMalloc((void**)&h->taps, sizeof(double) * n));
    if (nc) TF_CUDA(cudaMalloc((void**)&h->canal_taps, sizeof(double) * (size_t)h->taps_cap * nc));
    return 0;
}

// Queue n_steps on the handle's streams.  record != 0 stores the outlet taps
// of step s in slot s of the device tap buffer.
static int queue_steps(tf_handle h, int n_steps, int record) {
    const Geom& g = h->g;
    TF_CHECK(h->chan_ready, "tf_chan_step: call tf_chan_init first");
    TF_CHECK(n_steps >= 0, "tf_chan_step: n_steps < 0");
    if (record) TF_TRY(ensure_taps(h, n_steps));
    cudaStream_t s = h->stream;
    if (h->state_dirty) {
        ChanArgs a = chan_args(h, 0, g.ny);
        if (h->packed) k_init_mstate<true><<<tile_grid(g, g.ny), dim3(32, 8), 0, s>>>(a);
        else k_init_mstate<false><<<tile_grid(g, g.ny), dim3(32, 8), 0, s>>>(a);
        TF_LAUNCHED(h, s, "k_init_mstate");
        h->state_dirty = false;
    }
    if (!h->timing_open) {
        TF_CUDA(cudaEventRecord(h->t0, s));
        h->timing_open = true;
    }
    const bool any_points = h->pts[0].n || h->pts[1].n || h->pts[2].n || h->pts[3].n;
    const bool multi = h->nccl_comm != nullptr;
    for (int step = 0; step < n_steps; ++step) {
        // the grid written as "next Q" by the previous step is this step's Q
        double* t = h->Qcur; h->Qcur = h->Qnext; h->Qnext = t;
        if (h->wave == TF_DYNAMIC) {   // last step's u becomes u_old
            Raster<double> r = h->u; h->u = h->u2; h->u2 = r;
        }
        if (any_points) {
            PointsArgs p;
            for (int k = 0; k < 4; ++k) { p.n[k] = h->pts[k].n; p.idx[k] = h->pts[k].idx; p.val[k] = h->pts[k].val; }
            p.tmp = h->pts_tmp;
            p.canals_in_out = (record && h->canal_taps) ? h->canal_taps + (size_t)step * h->pts[TF_PT_CANALS_IN].n : nullptr;
            k_diversions<<<1, 32, 0, s>>>(p, h->Qcur, h->vol.p, h->dt);
            TF_LAUNCHED(h, s, "k_diversions");
            h->acc_launches++;
            if (multi && h->use_march) h->ghost_dirty = true;
            else if (multi) {   // a diverted cell on an edge row changes the neighbour's halo
                void* arr[1] = {h->Qcur};
                size_t es[1] = {sizeof(double)};
                TF_TRY(comm_exchange_rows(h, s, arr, es, 1));
            }
        }
        if (h->use_march) {
            if (h->ghost_dirty) TF_TRY(refresh_ghosts(h));
            TF_TRY(march_step(h));
        } else if (h->wave == TF_KINEMATIC) {
            if (!multi || g.ny < 32) {
                TF_TRY(launch_kinematic(h, s, 0, g.ny));
                if (multi) TF_TRY(exchange_planned(h, s, 1));
            } else {
                // edge bands first on the high-priority stream, their halo
                // rows travel while the interior is computed
                cudaStream_t e = h->stream_edge;
                TF_CUDA(cudaEventRecord(h->ev_main, s));
                TF_CUDA(cudaStreamWaitEvent(e, h->ev_main, 0));
                TF_TRY(launch_kinematic(h, e, 0, 8));
                TF_TRY(launch_kinematic(h, e, g.ny - 8, g.ny));
                TF_TRY(exchange_planned(h, e, 1));
                TF_CUDA(cudaEventRecord(h->ev_edge, e));
                TF_TRY(launch_kinematic(h, s, 8, g.ny - 8));
                TF_CUDA(cudaStreamWaitEvent(s, h->ev_edge, 0));
            }
        } else {
            TF_TRY(launch_phase_a(h, s, 0, g.ny));
            if (multi) TF_TRY(exchange_planned(h, s, 0));      // between the phases
            TF_TRY(launch_phase_b(h, s, 0, g.ny));
            if (multi) TF_TRY(exchange_planned(h, s, 1));      // at the end of the step
        }
        if (record && h->n_outlets) {
            ChanArgs a = chan_args(h, 0, g.ny);
            double* out = h->taps + (size_t)step * h->n_outlets * 4;
            TF_DISPATCH(h, (k_taps<P, L><<<(h->n_outlets + 63) / 64, 64, 0, s>>>(a, h->outlet_idx, h->n_outlets, out)));
            TF_LAUNCHED(h, s, "k_taps");
            h->acc_launches++;
        }
        h->steps_done++;         // the generation of the overbank byte follows the step count
    }
    h->acc_steps += n_steps;
    return 0;
}

extern "C" int tf_chan_step_async(tf_handle h, int n_steps) {
    TF_TRY(use_device(h));
    return queue_steps(h, n_steps, 0);
}

extern "C" int tf_chan_step(tf_handle h, int n_steps, double* taps_out, double* canals_in_out,
                            int64_t* bad_out) {
    TF_TRY(use_device(h));
    TF_TRY(queue_steps(h, n_steps, taps_out != nullptr || canals_in_out != nullptr));
    cudaStream_t s = h->stream;
    // Everything that comes back lands in one page-locked staging block first: copies into the
    // caller's pageable arrays would each be staged by the driver with a wait of its own; this
    // way the three copies are queued behind the kernels and the call waits once.
    const size_t n_taps = (taps_out && h->n_outlets && n_steps) ? (size_t)n_steps * h->n_outlets * 4 : 0;
    const int nc = h->pts[TF_PT_CANALS_IN].n;
    const size_t n_can = (canals_in_out && nc && n_steps) ? (size_t)n_steps * nc : 0;
    const size_t need = sizeof(double) * (n_taps + n_can) + 2 * sizeof(long long);
    if (need > h->back_cap) {
        if (h->back) cudaFreeHost(h->back);
        h->back = nullptr; h->back_cap = 0;
        const size_t cap = need < 65536 ? 65536 : need;
        TF_CUDA(cudaHostAlloc((void**)&h->back, cap, cudaHostAllocDefault));
        h->back_cap = cap;
    }
    long long* bad = reinterpret_cast<long long*>(h->back);
    double* b_taps = reinterpret_cast<double*>(h->back + 2 * sizeof(long long));
    double* b_can = b_taps + n_taps;
    if (n_taps) TF_CUDA(cudaMemcpyAsync(b_taps, h->taps, sizeof(double) * n_taps, cudaMemcpyDeviceToHost, s));
    if (n_can) TF_CUDA(cudaMemcpyAsync(b_can, h->canal_taps, sizeof(double) * n_can, cudaMemcpyDeviceToHost, s));
    TF_CUDA(cudaMemcpyAsync(bad, h->bad, 2 * sizeof(long long), cudaMemcpyDeviceToHost, s));
    TF_CUDA(cudaStreamSynchronize(s));
    if (n_taps) memcpy(taps_out, b_taps, sizeof(double) * n_taps);
    if (n_can) memcpy(canals_in_out, b_can, sizeof(double) * n_can);
    if (bad_out) {
        const long long none = 0x7fffffffffffffffLL;
        bad_out[0] = bad[0] == none ? -1 : bad[0];
        bad_out[1] = bad[1] == none ? -1 : bad[1];
    }
    return 0;
}

// Device pointer (cell (0,0), pitched) of a field: the state rasters as they
// stand, anything else derived into the scratch raster first.
int chan_derive_to_scratch(tf_handle h, int which, const double** src_out) {
    TF_CHECK(h->chan_ready, "tf_chan_get: call tf_chan_init first");
    TF_CHECK(which >= 0 && which < TF_F_COUNT, "tf_chan_get: unknown field %d", which);
    const Geom& g = h->g;
    switch (which) {
        case TF_F_VOL: *src_out = h->vol.p; return 0;
        case TF_F_D: *src_out = h->d.p; return 0;
        case TF_F_U: *src_out = h->u.p; return 0;
        case TF_F_Q: *src_out = h->Qcur; return 0;
        case TF_F_QNEXT: *src_out = h->Qnext; return 0;
        case TF_F_R:
            TF_CHECK(h->R_is_grid && h->R.p, "tf_chan_get: R is a scalar right now (%.17g)", h->R_scalar);
            *src_out = h->R.p; return 0;
        default: break;
    }
    TF_CHECK(!(which == TF_F_S_FREE && h->wave == TF_KINEMATIC),
             "tf_chan_get: S_free is not a device field for the kinematic wave (it stays the "
             "copy of S_bed taken at initialisation, channels_base.py:889)");
    if (!h->scratch64) TF_CUDA(cudaMalloc((void**)&h->scratch64, sizeof(double) * (size_t)g.n_alloc));
    if (h->state_dirty) {   // mutation flags not built yet: nothing has fired
        TF_CUDA(cudaMemsetAsync(h->mstate.base, 0, (size_t)g.n_alloc, h->stream));
    }
    ChanArgs a = chan_args(h, 0, g.ny);
    double* out = h->scratch64 + g.origin;
    dim3 blk(64, 4), grd((g.nx + 63) / 64, (g.ny + 3) / 4);
    TF_DISPATCH(h, (k_derive<P, L><<<grd, blk, 0, h->stream>>>(a, which, h->wave, out)));
    TF_CUDA(cudaGetLastError());
    *src_out = out;
    return 0;
}

extern "C" int tf_chan_get(tf_handle h, int which, double* host_out) {
    TF_TRY(use_device(h));
    TF_CHECK(host_out != nullptr, "tf_chan_get: host_out is NULL");
    const double* src = nullptr;
    TF_TRY(chan_derive_to_scratch(h, which, &src));
    return raster_download<double>(h, src, host_out);
}

extern "C" int tf_chan_minmax(tf_handle h, double* out6) {
    TF_TRY(use_device(h));
    TF_CHECK(h->chan_ready, "tf_chan_minmax: call tf_chan_init first");
    TF_CHECK(out6 != nullptr, "tf_chan_minmax: out6 is NULL");
    const Geom& g = h->g;
    if (h->partial_cap < 6 * g.ny) {
        if (h->partial) cudaFree(h->partial);
        h->partial_cap = 6 * g.ny;
        TF_CUDA(cudaMalloc((void**)&h->partial, sizeof(double) * h->partial_cap));
    }
    ChanArgs a = chan_args(h, 0, g.ny);
    k_minmax<<<g.ny, 256, 0, h->stream>>>(a, h->partial);
    TF_CUDA(cudaGetLastError());
    std::vector<double> part((size_t)6 * g.ny);
    TF_CUDA(cudaMemcpyAsync(part.data(), h->partial, sizeof(double) * part.size(), cudaMemcpyDeviceToHost, h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    for (int f = 0; f < 6; ++f) {
        double r = (f & 1) ? -INFINITY : INFINITY;
        bool nan_seen = false;
        for (int y = 0; y < g.ny; ++y) {
            const double v = part[(size_t)6 * y + f];
            if (v != v) nan_seen = true;
            else r = (f & 1) ? (v > r ? v : r) : (v < r ? v : r);
        }
        out6[f] = nan_seen ? NAN : r;
    }
    return 0;
}

static double layout_bytes_per_cell(tf_handle h) {
    // bytes the step kernels address per cell-update in the current layout
    double b = 0.0;
    const double shape = h->packed ? ((h->len_rows ? 4 : 4 + 4) + 1) : (8 + 8 + 8 + 8);   // w, ds, angle
    const double sr = h->packed ? 8 : 16;                            // S, roughness
    const double R = h->R_is_grid ? 8 : 0;
    const double dw = (h->packed && h->len_rows) ? 0 : 4;            // derived from the code
    if (h->wave == TF_KINEMATIC) {
        b = 8 /*Q*/ + 8 /*vol r*/ + 1 + 1 + 1 /*code, mask, mstate*/ + shape + dw
            + (h->kfac.p ? 8 : sr) + R + 32 /*vol, d, u, Qnext w*/;
    } else if (h->use_march) {
        // one fused pass: Q, vol, code, mask, mstate, w, angle index, S, n [, u_old] in;
        // vol, d, u, Qnext out
        b = 8 + 8 + 3 + shape + sr + R + 32 + (h->wave == TF_DYNAMIC ? 8 : 0);
    } else {
        const double A = 8 + 8 + 3 + shape + dw + R + 16;             // phase A
        double B = 8 /*d*/ + 2 + shape + dw + sr + 16;                // phase B: u, Qn
        if (h->wave == TF_DYNAMIC) B += 8 + 8 + 8 + 1;                // Q, vol, u_old, mask
        b = A + B;
    }
    return b;
}

extern "C" int tf_chan_timing(tf_handle h, double* out4) {
    TF_TRY(use_device(h));
    TF_CHECK(out4 != nullptr, "tf_chan_timing: out4 is NULL");
    if (h->timing_open) {
        TF_CUDA(cudaEventRecord(h->t1, h->stream));
        TF_CUDA(cudaEventSynchronize(h->t1));
        float ms = 0.f;
        TF_CUDA(cudaEventElapsedTime(&ms, h->t0, h->t1));
        h->acc_ms += ms;
        h->timing_open = false;
    }
    out4[0] = h->acc_ms;
    out4[1] = (double)h->acc_steps;
    out4[2] = (double)h->acc_launches;
    out4[3] = h->chan_ready ? layout_bytes_per_cell(h) : 0.0;
    h->acc_ms = 0.0; h->acc_steps = 0; h->acc_launches = 0;
    return 0;
}

int chan_free(tf_handle h) {
    raster_free(h->vol); raster_free(h->volB); raster_free(h->d); raster_free(h->u); raster_free(h->u2);
    raster_free(h->Qa); raster_free(h->Qb); raster_free(h->R); raster_free(h->R1);
    for (int k = 0; k < TF_R_COUNT; ++k) { raster_free(h->forc[k]); h->forc_is_grid[k] = false; }
    raster_free(h->mstate);
    raster_free(h->w64); raster_free(h->tan64); raster_free(h->cos64);
    raster_free(h->S64); raster_free(h->n64); raster_free(h->ds64); raster_free(h->kfac);
    raster_free(h->w32); raster_free(h->S32); raster_free(h->n32); raster_free(h->ds32);
    raster_free(h->aidx);
    void* ptrs[] = {h->lut, h->len_rows, h->da_rows, h->outlet_idx, h->taps, h->canal_taps,
                    h->partial, h->scratch64, h->pts_tmp};
    for (void* p : ptrs) if (p) cudaFree(p);
    h->lut = nullptr; h->len_rows = nullptr; h->da_rows = nullptr; h->outlet_idx = nullptr;
    h->taps = nullptr; h->canal_taps = nullptr; h->partial = nullptr; h->scratch64 = nullptr;
    h->pts_tmp = nullptr;
    for (int k = 0; k < 4; ++k) {
        if (h->pts[k].idx) cudaFree(h->pts[k].idx);
        if (h->pts[k].val) cudaFree(h->pts[k].val);
        h->pts[k] = PointSet();
    }
    if (h->back) cudaFreeHost(h->back);
    h->back = nullptr; h->back_cap = 0;
    h->chan_ready = false;
    return 0;
}
