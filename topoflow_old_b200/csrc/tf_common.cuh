// tf_common.cuh -- handle layout, device-grid geometry and error plumbing
// shared by the D8, routing and communication translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>

#include "../../include/topoflow_b200.h"

// ---------------------------------------------------------------------------
// Device grid geometry.
//
// Every device raster shares one layout: (ny + 2 * TF_APRON) rows of `pitch`
// elements, pitch = nx rounded up to 32, preceded and followed by TF_GUARD
// guard elements.  Rows -1 and ny are apron rows (the neighbour strip's edge
// row, or zeros at a global border), so a kernel may address (y-1 .. y+1,
// x-1 .. x+1) of any owned cell without bounds tests: x = -1 lands in the
// previous row's padding (or, when pitch == nx, on the previous row's last
// cell, which is a border cell with flow code 0 and never contributes).  The
// second apron rows (-2 and ny + 1) exist for the fused diffusive / dynamic
// step, which recomputes the depth of the first apron row and therefore reads
// the discharge one row further out.  All row starts are 32-element aligned,
// so 128-bit loads are always legal; the guards are wide enough for a 64-cell
// window of the marching kernels to overhang a row end.
// ---------------------------------------------------------------------------
struct Geom {
    int nx, ny;          // owned cells
    int pitch;           // elements per row
    int64_t n_alloc;     // elements per raster incl. aprons and guards
    int64_t origin;      // element offset of cell (0, 0)
    int row_offset;      // first owned row in the global grid
    int global_ny;       // rows of the global grid
    int has_north, has_south;   // a neighbour strip exists on that side
};

static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

#define TF_GUARD 128
#define TF_APRON 2      // apron rows above and below the owned rows
#define TF_LEN_ROW 16   // floats per row of the length table (tf_handle_s::len_rows)
#define TF_STR2(x) #x
#define TF_STR(x) TF_STR2(x)

// ---------------------------------------------------------------------------
// Error handling
// ---------------------------------------------------------------------------
void tf_set_error(const char* fmt, ...);

#define TF_CUDA(call)                                                          \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) {                                               \
            tf_set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,  \
                         cudaGetErrorString(e_));                              \
            return 1;                                                          \
        }                                                                      \
    } while (0)

#define TF_CHECK(cond, ...)                                                    \
    do {                                                                       \
        if (!(cond)) {                                                         \
            tf_set_error(__VA_ARGS__);                                         \
            return 1;                                                          \
        }                                                                      \
    } while (0)

// after a kernel launch: always catch launch errors; with TF_B200_DEBUG_SYNC=1
// also wait for the kernel so a device fault is reported against its name
#define TF_LAUNCHED(h, stream, name)                                           \
    do {                                                                       \
        cudaError_t e_ = cudaGetLastError();                                   \
        if (e_ == cudaSuccess && (h)->debug_sync) e_ = cudaStreamSynchronize(stream); \
        if (e_ != cudaSuccess) {                                               \
            tf_set_error("kernel %s: %s", name, cudaGetErrorString(e_));       \
            return 1;                                                          \
        }                                                                      \
    } while (0)

#define TF_TRY(call)                                                           \
    do {                                                                       \
        int r_ = (call);                                                       \
        if (r_) return r_;                                                     \
    } while (0)

// ---------------------------------------------------------------------------
// Typed device raster
// ---------------------------------------------------------------------------
template <typename T>
struct Raster {
    T* base = nullptr;   // allocation start
    T* p = nullptr;      // pointer to cell (0, 0)
};

struct PointSet {
    int n = 0;
    int64_t* idx = nullptr;   // device: element offsets relative to cell (0,0)
    double* val = nullptr;    // device
};

struct HostParam {
    Raster<double> stage;       // the grid as it was handed over (device, float64); unset when scalar
    double scalar = 0.0;
    int is_grid = 0;            // grid-typed in the reference
    bool set = false;
};

struct NcclApi;   // tf_comm.cu

struct tf_handle_s {
    int device = 0;
    Geom g{};
    cudaStream_t stream = nullptr;        // main stream
    cudaStream_t stream_edge = nullptr;   // boundary rows + halo exchange
    cudaEvent_t ev_edge = nullptr, ev_main = nullptr;

    // ---- D8 -------------------------------------------------------------
    Raster<float> dem;            // DEM, float32
    Raster<int16_t> raw, raw2;    // working codes (Jacobi pair for link_flats)
    Raster<uint8_t> code;         // final D8 codes
    Raster<uint8_t> inmask;       // bit k: child in the k-th add direction
    Raster<int32_t> parent;       // parent_ID_grid (global calendar IDs)
    Raster<float> ds0, dw;        // flow length / width (float32)
    Raster<float> area;           // contributing area
    Raster<int32_t> pending;      // area: children still to finish
    float *dx = nullptr, *dy = nullptr, *dd = nullptr;   // per-row, float32
    bool have_d8 = false, have_area = false;
    long long area_rounds = 0;            // strips: exchange rounds the last tf_d8_area took
    int lr_periodic = 0, tb_periodic = 0;
    float row0_len[3] = {0.f, 0.f, 0.f};  // dx, dy, dd of GLOBAL row 0 (tf_d8_set_row0_sizes)
    bool have_row0_len = false;
    unsigned long long* counters = nullptr;   // device scratch counters [16]

    // ---- routing ----------------------------------------------------------
    HostParam hp[TF_P_COUNT];
    // bank angles handed over as angles (tf_chan_set_angle / tf_chan_set_angle_table)
    std::vector<double> angle_values, angle_tan, angle_cos;
    int angle_table_n = -1;               // > 0: the table is complete, aidx indexes it (+1)
    bool angle_scalar = false, angle_is_grid = false;
    int wave = 0, law = 0, packed = 0;
    int mut_width = 0, mut_angle = 0;
    double dt = 0.0, rho = 1000.0;
    long long steps_done = 0;             // since tf_chan_init
    bool state_dirty = true;              // mstate must be rebuilt from d
    double* da_rows = nullptr;            // device [ny + 2 * TF_APRON] (index y + TF_APRON)
    bool chan_ready = false;
    // state
    Raster<double> vol, volB, d, u, u2, Qa, Qb, R;   // volB: ping-pong partner (fused marching step)
    bool use_march = false;               // marching kernels (tf_chan_march.cuh) instead of the tile kernels
    int seg_rows = 64;                    // rows per warp task of the marching kernels (TF_B200_SEG_ROWS)
    bool ghost_dirty = true;              // apron copies of vol / mstate / Q / u must be refreshed from the neighbours
    double *Qcur = nullptr, *Qnext = nullptr;   // -> Qa.p / Qb.p
    bool R_is_grid = false;
    double R_scalar = 0.0;
    Raster<double> forc[TF_R_COUNT];      // fluxes of update_R handed over as grids
    double forc_scalar[TF_R_COUNT] = {0, 0, 0, 0, 0, 0};
    bool forc_is_grid[TF_R_COUNT] = {false, false, false, false, false, false};
    // device-resident providers (tf_forcing.cu)
    std::vector<Raster<double>> grids;    // tf_grid_new
    const double* fdev_value[TF_R_COUNT] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    const double* fdev_rate[TF_R_COUNT] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    double fdev_value_s[TF_R_COUNT] = {0, 0, 0, 0, 0, 0}, fdev_rate_s[TF_R_COUNT] = {0, 0, 0, 0, 0, 0};
    Raster<double> R1;                    // dR/dt while a provider is being interpolated: R = R1 * t + R
    bool R_rate_on = false;
    Raster<uint8_t> mstate;               // bit0 mutated, bit1 pending
    // parameters, generic layout (float64)
    Raster<double> w64, tan64, cos64, S64, n64, ds64;
    Raster<double> kfac;                  // sqrt(S_bed)/n for the kinematic Manning step
    // parameters, packed layout (float32 + angle table)
    Raster<float> w32, S32, n32, ds32;
    Raster<uint8_t> aidx;
    double* lut = nullptr;                // device [3*256]: tan, cos, 1 / (2 tan)
    bool rcp_ok = true;                   // every reciprocal the marching kernels divide by is safe
    float* len_rows = nullptr;            // device [TF_LEN_ROW * (ny + 2 * TF_APRON)], row y at TF_LEN_ROW * (y + TF_APRON):
                                          // ds {diag, E/W, N/S, -}, dw {diag, E/W, N/S, -} float32, then 1/ds {diag, E/W, N/S, -} float64
    // outlets, diversions
    int n_outlets = 0;
    int64_t* outlet_idx = nullptr;        // device
    std::vector<int64_t> outlet_idx_h;
    PointSet pts[4];
    double* pts_tmp = nullptr;            // staging for the diversion kernel
    int pts_tmp_cap = 0;
    double* taps = nullptr;               // device ring [cap][n_outlets][4]
    double* canal_taps = nullptr;         // device [cap][n_canals]
    int taps_cap = 0;
    long long* bad = nullptr;             // device [2]
    char* back = nullptr;                 // page-locked staging of what tf_chan_step copies back
    size_t back_cap = 0;
    double* partial = nullptr;            // device scratch for reductions
    int partial_cap = 0;
    double* scratch64 = nullptr;          // one raster for derived fields
    // output frames (tf_output.cu)
    float* frame32 = nullptr;             // device [ny*nx] dense float32 frame
    cudaStream_t stream_io = nullptr;     // D2H copies of frames
    cudaEvent_t ev_frame_ready = nullptr, ev_frame_done = nullptr;
    bool frame_pending = false;
    // timing
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    bool timing_open = false;
    double acc_ms = 0.0;
    long long acc_steps = 0, acc_launches = 0;

    // ---- communication ------------------------------------------------------
    void* nccl_comm = nullptr;
    int rank = 0, nranks = 1;
    int pf_rows = -1;                     // L2 look-ahead rows; -1 = from the grid width (TF_B200_PF_ROWS)
    bool debug_sync = false;              // TF_B200_DEBUG_SYNC=1: sync + check after every launch
};

// helpers implemented in tf_api.cu
template <typename T> int raster_alloc(tf_handle h, Raster<T>& r, int fill_byte);
template <typename T> void raster_free(Raster<T>& r);
template <typename T> int raster_upload(tf_handle h, Raster<T>& r, const T* host);
template <typename T> int raster_download(tf_handle h, const T* dev_p, T* host);
int use_device(tf_handle h);

// implemented per TU
int d8_free(tf_handle h);
int chan_free(tf_handle h);
int comm_free(tf_handle h);
int output_free(tf_handle h);
int forcing_free(tf_handle h);
int chan_derive_to_scratch(tf_handle h, int which, const double** src_out);
int comm_exchange_rows(tf_handle h, cudaStream_t s, void* const* arrays,
                       const size_t* elem_sizes, int n_arrays, const int* depths = nullptr);
int comm_exchange_bytes(tf_handle h, cudaStream_t s, const void* send_north, void* recv_north,
                        const void* send_south, void* recv_south, size_t bytes);
int comm_allreduce_u64(tf_handle h, cudaStream_t s, unsigned long long* dev, int n);
int comm_allgather_rows(tf_handle h, cudaStream_t s, const void* own, void* full, size_t row_bytes,
                        const int* first_row, const int* n_rows);
