// tf_chan_host.cuh -- host helpers shared by the routing translation units.
#pragma once
#include "tf_chan_march.cuh"
#include "tf_common.cuh"

// ---------------------------------------------------------------------------
static inline ChanArgs chan_args(tf_handle h, int y_begin, int y_end) {
    ChanArgs a;
    memset(&a, 0, sizeof(a));
    const Geom& g = h->g;
    a.nx = g.nx; a.ny = g.ny; a.pitch = g.pitch;
    a.y_begin = y_begin; a.y_end = y_end;
    if (h->pf_rows >= 0) a.pf_rows = h->pf_rows;
    else a.pf_rows = 8;        // the next row of tiles: best at 8192^2 and 16384^2 (profiles/sweep_r01.txt)
    a.row_offset = g.row_offset; a.global_ny = g.global_ny;
    a.dt = h->dt; a.rho = h->rho;
    a.Q = h->Qcur; a.Qn = h->Qnext;
    a.vol = h->vol.p; a.d = h->d.p; a.u = h->u.p; a.u_old = h->u2.p;
    a.steps_done = h->steps_done;
    a.Rgrid = h->R_is_grid ? h->R.p : nullptr;
    a.Rs = h->R_scalar;
    a.Rrate = (h->R_is_grid && h->R_rate_on) ? h->R1.p : nullptr;
    a.t_now = (double)h->steps_done * h->dt;
    a.da_rows = h->da_rows + TF_APRON;
    a.code = h->code.p; a.inmask = h->inmask.p; a.dw = h->dw.p;
    a.mstate = h->mstate.p;
    a.mut_width = h->mut_width; a.mut_angle = h->mut_angle;
    a.w64 = h->w64.p; a.tan64 = h->tan64.p; a.cos64 = h->cos64.p;
    a.S64 = h->S64.p; a.n64 = h->n64.p; a.ds64 = h->ds64.p;
    a.w32 = h->w32.p; a.S32 = h->S32.p; a.n32 = h->n32.p; a.ds32 = h->ds32.p;
    a.aidx = h->aidx.p; a.lut = h->lut;
    a.kfac = h->kfac.p;
    a.len_rows = h->len_rows ? h->len_rows + TF_LEN_ROW * TF_APRON : nullptr;
    a.bad = h->bad;
    // marching kernels: the fused waves write the volume into the partner buffer
    a.march = h->use_march ? 1 : 0;
    a.vol_out = (h->use_march && h->wave != TF_KINEMATIC) ? h->volB.p : h->vol.p;
    a.n_bands = 0; a.seg_rows = h->seg_rows;
    a.has_north = g.has_north; a.has_south = g.has_south;
    a.rcp_ok = h->rcp_ok ? 1 : 0;
    return a;
}

static inline dim3 tile_grid(const Geom& g, int rows) {
    return dim3((g.pitch / CPT + 31) / 32, (rows + 7) / 8);
}

// PACKED x LAW dispatch
#define TF_DISPATCH(h, EXPR)                                              \
    do {                                                                  \
        if ((h)->packed) {                                                \
            if ((h)->law == 0) { constexpr bool P = true; constexpr int L = 0; EXPR; }  \
            else { constexpr bool P = true; constexpr int L = 1; EXPR; }  \
        } else {                                                          \
            if ((h)->law == 0) { constexpr bool P = false; constexpr int L = 0; EXPR; } \
            else { constexpr bool P = false; constexpr int L = 1; EXPR; } \
        }                                                                 \
    } while (0)

