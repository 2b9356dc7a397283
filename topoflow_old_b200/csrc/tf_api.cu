// tf_api.cu -- handle lifetime, error reporting and raster transfer helpers.
#include <stdarg.h>
#include <stdlib.h>

#include "tf_common.cuh"

static thread_local char g_err[1024] = "";

void tf_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* tf_last_error(void) { return g_err; }

extern "C" const char* tf_version(void) {
    return "topoflow_b200 0.2 (sm_100a, CUDA " TF_STR(CUDART_VERSION) ")";
}

extern "C" int tf_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        tf_set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return -1;
    }
    return n;
}

int use_device(tf_handle h) {
    TF_CHECK(h != nullptr, "null handle");
    TF_CUDA(cudaSetDevice(h->device));
    return 0;
}

// ---------------------------------------------------------------------------
template <typename T>
int raster_alloc(tf_handle h, Raster<T>& r, int fill_byte) {
    if (r.base) return 0;
    TF_CUDA(cudaMalloc((void**)&r.base, sizeof(T) * (size_t)h->g.n_alloc));
    TF_CUDA(cudaMemsetAsync(r.base, fill_byte, sizeof(T) * (size_t)h->g.n_alloc,
                            h->stream));
    r.p = r.base + h->g.origin;
    return 0;
}

template <typename T>
void raster_free(Raster<T>& r) {
    if (r.base) cudaFree(r.base);
    r.base = r.p = nullptr;
}

template <typename T>
int raster_upload(tf_handle h, Raster<T>& r, const T* host) {
    const Geom& g = h->g;
    TF_CUDA(cudaMemcpy2DAsync(r.p, sizeof(T) * g.pitch, host, sizeof(T) * g.nx,
                              sizeof(T) * g.nx, g.ny, cudaMemcpyHostToDevice,
                              h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
}

template <typename T>
int raster_download(tf_handle h, const T* dev_p, T* host) {
    const Geom& g = h->g;
    TF_CUDA(cudaMemcpy2DAsync(host, sizeof(T) * g.nx, dev_p, sizeof(T) * g.pitch,
                              sizeof(T) * g.nx, g.ny, cudaMemcpyDeviceToHost,
                              h->stream));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
}

#define TF_INST(T)                                                    \
    template int raster_alloc<T>(tf_handle, Raster<T>&, int);         \
    template void raster_free<T>(Raster<T>&);                         \
    template int raster_upload<T>(tf_handle, Raster<T>&, const T*);   \
    template int raster_download<T>(tf_handle, const T*, T*);
TF_INST(float)
TF_INST(double)
TF_INST(uint8_t)
TF_INST(int16_t)
TF_INST(int32_t)

// ---------------------------------------------------------------------------
extern "C" int tf_create(int nx, int ny, int device, unsigned flags,
                         tf_handle* out) {
    TF_CHECK(out != nullptr, "tf_create: out is NULL");
    *out = nullptr;
    TF_CHECK(nx >= 3 && ny >= 3, "tf_create: grid must be at least 3 x 3 (got %d x %d)", nx, ny);
    TF_CHECK((int64_t)nx * ny < ((int64_t)1 << 31),
             "tf_create: nx*ny = %lld does not fit the reference's int32 cell IDs",
             (long long)nx * ny);
    int ndev = tf_device_count();
    TF_CHECK(ndev > 0, "tf_create: no CUDA device available (%s)", g_err);
    TF_CHECK(device >= 0 && device < ndev, "tf_create: device %d out of range [0, %d)", device, ndev);
    (void)flags;
    tf_handle h = new tf_handle_s();
    h->device = device;
    { const char* e = getenv("TF_B200_DEBUG_SYNC"); h->debug_sync = e && e[0] == '1'; }
    { const char* e = getenv("TF_B200_PF_ROWS"); if (e && e[0]) h->pf_rows = atoi(e); }
    Geom& g = h->g;
    g.nx = nx;
    g.ny = ny;
    g.pitch = round_up(nx, 32);
    g.n_alloc = (int64_t)(ny + 2 * TF_APRON) * g.pitch + 2 * TF_GUARD;
    g.origin = TF_GUARD + (int64_t)TF_APRON * g.pitch;
    g.row_offset = 0;
    g.global_ny = ny;
    g.has_north = g.has_south = 0;
    if (cudaSetDevice(device) != cudaSuccess) {
        tf_set_error("tf_create: cudaSetDevice(%d) failed", device);
        delete h;
        return 1;
    }
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    cudaError_t e = cudaStreamCreateWithPriority(&h->stream, cudaStreamNonBlocking, lo);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&h->stream_edge, cudaStreamNonBlocking, hi);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_edge, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_main, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreate(&h->t0);
    if (e == cudaSuccess) e = cudaEventCreate(&h->t1);
    if (e == cudaSuccess) e = cudaMalloc((void**)&h->counters, 16 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc((void**)&h->bad, 2 * sizeof(long long));
    if (e != cudaSuccess) {
        tf_set_error("tf_create: %s", cudaGetErrorString(e));
        tf_destroy(h);
        return 1;
    }
    *out = h;
    return 0;
}

extern "C" int tf_set_strip(tf_handle h, int row_offset, int global_ny,
                            int has_north, int has_south) {
    TF_TRY(use_device(h));
    TF_CHECK(!h->have_d8, "tf_set_strip: call before tf_d8_build");
    TF_CHECK(row_offset >= 0 && row_offset + h->g.ny <= global_ny,
             "tf_set_strip: rows [%d, %d) outside a grid of %d rows",
             row_offset, row_offset + h->g.ny, global_ny);
    // A global grid of >= 2^31 cells cannot carry the reference's int32 calendar
    // IDs (d8_base.py:675); routing does not need them, so only the parent-ID
    // product becomes unavailable (tf_d8_get refuses it).
    h->g.row_offset = row_offset;
    h->g.global_ny = global_ny;
    h->g.has_north = has_north != 0;
    h->g.has_south = has_south != 0;
    return 0;
}

extern "C" int tf_destroy(tf_handle h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    comm_free(h);
    output_free(h);
    forcing_free(h);
    chan_free(h);
    d8_free(h);
    if (h->counters) cudaFree(h->counters);
    if (h->bad) cudaFree(h->bad);
    if (h->t0) cudaEventDestroy(h->t0);
    if (h->t1) cudaEventDestroy(h->t1);
    if (h->ev_edge) cudaEventDestroy(h->ev_edge);
    if (h->ev_main) cudaEventDestroy(h->ev_main);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->stream_edge) cudaStreamDestroy(h->stream_edge);
    delete h;
    return 0;
}

extern "C" int tf_sync(tf_handle h) {
    TF_TRY(use_device(h));
    TF_CUDA(cudaStreamSynchronize(h->stream_edge));
    TF_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
}
