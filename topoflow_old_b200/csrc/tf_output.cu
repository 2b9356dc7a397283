// tf_output.cu -- the output side of the routing path: grid frames as the
// reference writes them to its .rts stacks (utils/rts_files.py:313-396:
// float32, byte-swapped when the RTI byte order differs from the host's),
// converted on the device and copied back on a stream of their own so the
// next routing steps overlap the transfer.
#include "tf_common.cuh"

// np.float32(double) rounds to nearest even; cvt.rn.f32.f64 is the same rounding.
static __global__ void __launch_bounds__(256)
k_frame_f32(const double* __restrict__ src, int nx, int ny, int pitch, int swap,
            uint32_t* __restrict__ dst) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= nx || y >= ny) return;
    const float v = __double2float_rn(__ldcs(src + (int64_t)y * pitch + x));
    uint32_t bits = __float_as_uint(v);
    if (swap) bits = __byte_perm(bits, 0u, 0x0123);
    __stcs(dst + (int64_t)y * nx + x, bits);
}

static int frame_setup(tf_handle h) {
    const Geom& g = h->g;
    if (!h->frame32) TF_CUDA(cudaMalloc((void**)&h->frame32, sizeof(float) * (size_t)g.nx * g.ny));
    if (!h->stream_io) TF_CUDA(cudaStreamCreateWithFlags(&h->stream_io, cudaStreamNonBlocking));
    if (!h->ev_frame_ready) TF_CUDA(cudaEventCreateWithFlags(&h->ev_frame_ready, cudaEventDisableTiming));
    if (!h->ev_frame_done) TF_CUDA(cudaEventCreateWithFlags(&h->ev_frame_done, cudaEventDisableTiming));
    return 0;
}

extern "C" int tf_chan_frame(tf_handle h, int which, int swap_endian, float* host_out,
                             int async) {
    TF_TRY(use_device(h));
    TF_CHECK(host_out != nullptr, "tf_chan_frame: host_out is NULL");
    TF_TRY(frame_setup(h));
    const Geom& g = h->g;
    const double* src = nullptr;
    TF_TRY(chan_derive_to_scratch(h, which, &src));
    // the one device frame buffer is free again once the previous copy has left it
    if (h->frame_pending) TF_CUDA(cudaStreamWaitEvent(h->stream, h->ev_frame_done, 0));
    dim3 blk(64, 4), grd((g.nx + 63) / 64, (g.ny + 3) / 4);
    k_frame_f32<<<grd, blk, 0, h->stream>>>(src, g.nx, g.ny, g.pitch, swap_endian ? 1 : 0,
                                            (uint32_t*)h->frame32);
    TF_LAUNCHED(h, h->stream, "k_frame_f32");
    TF_CUDA(cudaEventRecord(h->ev_frame_ready, h->stream));
    TF_CUDA(cudaStreamWaitEvent(h->stream_io, h->ev_frame_ready, 0));
    TF_CUDA(cudaMemcpyAsync(host_out, h->frame32, sizeof(float) * (size_t)g.nx * g.ny,
                            cudaMemcpyDeviceToHost, h->stream_io));
    TF_CUDA(cudaEventRecord(h->ev_frame_done, h->stream_io));
    h->frame_pending = true;
    if (!async) return tf_chan_frame_wait(h);
    return 0;
}

extern "C" int tf_chan_frame_wait(tf_handle h) {
    TF_TRY(use_device(h));
    if (h->frame_pending) {
        TF_CUDA(cudaEventSynchronize(h->ev_frame_done));
        h->frame_pending = false;
    }
    return 0;
}

extern "C" int tf_host_alloc(size_t bytes, void** out) {
    TF_CHECK(out != nullptr, "tf_host_alloc: out is NULL");
    *out = nullptr;
    TF_CUDA(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return 0;
}

extern "C" int tf_host_free(void* p) {
    if (p) TF_CUDA(cudaFreeHost(p));
    return 0;
}

int output_free(tf_handle h) {
    if (h->frame_pending && h->ev_frame_done) cudaEventSynchronize(h->ev_frame_done);
    h->frame_pending = false;
    if (h->frame32) cudaFree(h->frame32);
    if (h->stream_io) cudaStreamDestroy(h->stream_io);
    if (h->ev_frame_ready) cudaEventDestroy(h->ev_frame_ready);
    if (h->ev_frame_done) cudaEventDestroy(h->ev_frame_done);
    h->frame32 = nullptr; h->stream_io = nullptr;
    h->ev_frame_ready = nullptr; h->ev_frame_done = nullptr;
    return 0;
}
