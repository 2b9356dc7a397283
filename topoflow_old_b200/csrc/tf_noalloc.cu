// tf_noalloc.cu -- runoff-volume integral with no allocation at all.
//
// Some callers of this library must not allocate device memory: code that
// runs between cudaStreamBeginCapture / EndCapture, plug-ins hosted by an
// application whose memory all comes from its own pool (a cudaMalloc behind
// its back would break its accounting and can stall every stream of the
// process), and tight time loops that cannot afford the device-wide
// synchronisation a cudaMalloc / cudaFree may imply.  tf_chan_set_R computes
// sum((R*da)*dt) with per-row partial sums kept in a cudaMalloc'ed scratch
// buffer owned by the handle; this file offers the same sum with the partials
// in a fixed array that is part of the library image instead.
//
// The array costs TF_NOALLOC_MAX_ROWS * 8 bytes (512 KiB) on every GPU that
// loads the library, for the life of the process.
#include "tf_common.cuh"

#define TF_NOALLOC_MAX_ROWS 65536
#define TF_NOALLOC_MAX_DEVICES 64

// Scratch: one partial sum per grid row.  This name is only a device symbol;
// host code must never pass it to a kernel -- it goes through
// cudaGetSymbolAddress (below).
static __device__ __align__(256) double g_row_partials[TF_NOALLOC_MAX_ROWS];

// per-device address of g_row_partials, resolved on first use on that device
static double* g_partials_addr[TF_NOALLOC_MAX_DEVICES] = {nullptr};

// Row y of (R*da)*dt summed by one block into partials[y] (same arithmetic
// and the same in-block order as k_sum_R, update_R_integral,
// channels_base.py:1151-1160).  Every slot 0 .. ny-1 is overwritten, so the
// previous call's contents never leak into this one and nothing needs
// clearing first.
static __global__ void __launch_bounds__(256)
k_sum_R_rows_static(const double* __restrict__ R, int nx, long long pitch,
                    const double* __restrict__ da_rows, double da, double dt,
                    double* partials) {
    __shared__ double sh[256];
    const int y = blockIdx.x;
    const double a = da_rows ? da_rows[y] : da;
    double acc = 0.0;
    for (int x = threadIdx.x; x < nx; x += blockDim.x)
        acc += (R[(long long)y * pitch + x] * a) * dt;
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partials[y] = sh[0];
}

// Rows added north to south by one thread: the order tf_chan_set_R uses on
// the host, so both entry points return the same bits.
static __global__ void k_sum_rows_static(const double* partials, int ny, double* out) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    double s = 0.0;
    for (int y = 0; y < ny; ++y) s += partials[y];
    *out = s;
}

extern "C" int tf_sum_R_noalloc(const double* R_dev, int nx, int ny, long long pitch,
                                const double* da_rows_dev, double da, double dt,
                                double* sum_out_dev, void* stream) {
    TF_CHECK(R_dev != nullptr && sum_out_dev != nullptr, "tf_sum_R_noalloc: NULL device pointer");
    TF_CHECK(nx > 0 && ny > 0 && pitch >= nx, "tf_sum_R_noalloc: bad shape %d x %d, pitch %lld", ny, nx, pitch);
    TF_CHECK(ny <= TF_NOALLOC_MAX_ROWS,
             "tf_sum_R_noalloc: %d rows exceed the built-in scratch of %d rows; use tf_chan_set_R, "
             "which allocates what it needs", ny, TF_NOALLOC_MAX_ROWS);
    int dev = -1;
    TF_CUDA(cudaGetDevice(&dev));
    TF_CHECK(dev >= 0 && dev < TF_NOALLOC_MAX_DEVICES, "tf_sum_R_noalloc: device %d not supported", dev);
    if (!g_partials_addr[dev]) {
        // set-up, once per device, with that device current: ask the runtime
        // where this device keeps its copy of the array
        double* p = nullptr;
        TF_CUDA(cudaGetSymbolAddress((void**)&p, g_row_partials));
        TF_CHECK(p != nullptr, "tf_sum_R_noalloc: cudaGetSymbolAddress returned NULL");
        g_partials_addr[dev] = p;
    }
    double* partials = g_partials_addr[dev];
    cudaStream_t s = (cudaStream_t)stream;
    k_sum_R_rows_static<<<ny, 256, 0, s>>>(R_dev, nx, pitch, da_rows_dev, da, dt, partials);
    TF_CUDA(cudaGetLastError());
    k_sum_rows_static<<<1, 32, 0, s>>>(partials, ny, sum_out_dev);
    TF_CUDA(cudaGetLastError());
    return 0;
}
