// tf_comm.cu -- halo rows between row strips over NCCL send/recv.
//
// The reference is a single process and has no communication of any kind.
// Here a grid may be cut into row strips, one per GPU / process; the only
// data a strip needs from a neighbour is that neighbour's edge row of a few
// rasters, which lands in the strip's apron row (tf_common.cuh).  NCCL is
// loaded with dlopen so that a single-GPU user needs no NCCL at all.
#include <dlfcn.h>

#include "tf_common.cuh"

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclInt8 = 0, ncclUint64 = 5 };
enum { ncclSum = 0 };

struct NcclFns {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclFns g_nccl;

static int nccl_load() {
    if (g_nccl.lib) return 0;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* lib = nullptr;
    for (const char* n : names) {
        lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    TF_CHECK(lib != nullptr, "NCCL not found (dlopen libnccl.so.2): %s", dlerror());
#define TF_SYM(field, name)                                                   \
    *(void**)(&g_nccl.field) = dlsym(lib, name);                              \
    TF_CHECK(g_nccl.field != nullptr, "NCCL symbol %s missing", name)
    TF_SYM(GetUniqueId, "ncclGetUniqueId");
    TF_SYM(CommInitRank, "ncclCommInitRank");
    TF_SYM(CommDestroy, "ncclCommDestroy");
    TF_SYM(GroupStart, "ncclGroupStart");
    TF_SYM(GroupEnd, "ncclGroupEnd");
    TF_SYM(Send, "ncclSend");
    TF_SYM(Recv, "ncclRecv");
    TF_SYM(AllReduce, "ncclAllReduce");
    TF_SYM(GetErrorString, "ncclGetErrorString");
#undef TF_SYM
    g_nccl.lib = lib;
    return 0;
}

#define TF_NCCL(call)                                                         \
    do {                                                                      \
        ncclResult_t r_ = (call);                                             \
        if (r_ != 0) {                                                        \
            tf_set_error("%s failed: %s", #call, g_nccl.GetErrorString(r_));  \
            return 1;                                                         \
        }                                                                     \
    } while (0)

extern "C" int tf_comm_unique_id(uint8_t* id128) {
    TF_CHECK(id128 != nullptr, "tf_comm_unique_id: NULL buffer");
    TF_TRY(nccl_load());
    ncclUniqueId id;
    TF_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id128, id.internal, 128);
    return 0;
}

extern "C" int tf_comm_init(tf_handle h, int rank, int nranks, const uint8_t* id128) {
    TF_TRY(use_device(h));
    TF_CHECK(id128 != nullptr, "tf_comm_init: NULL id");
    TF_CHECK(nranks >= 1 && rank >= 0 && rank < nranks, "tf_comm_init: rank %d of %d", rank, nranks);
    TF_CHECK(h->nccl_comm == nullptr, "tf_comm_init: communicator already set");
    h->rank = rank;
    h->nranks = nranks;
    if (nranks == 1) return 0;
    TF_CHECK(h->g.has_north == (rank > 0) && h->g.has_south == (rank < nranks - 1),
             "tf_comm_init: tf_set_strip neighbours do not match rank %d of %d (strips are "
             "stacked north to south in rank order)", rank, nranks);
    TF_TRY(nccl_load());
    ncclUniqueId id;
    memcpy(id.internal, id128, 128);
    ncclComm_t comm = nullptr;
    TF_NCCL(g_nccl.CommInitRank(&comm, nranks, id, rank));
    h->nccl_comm = comm;
    return 0;
}

// For each raster: my first row -> north neighbour's south apron, my last row
// -> south neighbour's north apron, and the mirror receives.  One NCCL group,
// queued on stream s.
// depths[k] (default 1, at most TF_APRON) = how many edge rows of raster k go
// each way: the fused marching step reads the discharge two rows out.
int comm_exchange_rows(tf_handle h, cudaStream_t s, void* const* arrays,
                       const size_t* elem_sizes, int n_arrays, const int* depths) {
    if (!h->nccl_comm) return 0;
    const Geom& g = h->g;
    ncclComm_t comm = (ncclComm_t)h->nccl_comm;
    TF_NCCL(g_nccl.GroupStart());
    for (int k = 0; k < n_arrays; ++k) {
        char* p = (char*)arrays[k];
        const size_t row = (size_t)g.pitch * elem_sizes[k];
        const int d = depths ? depths[k] : 1;
        TF_CHECK(d >= 1 && d <= TF_APRON && d <= g.ny, "comm_exchange_rows: depth %d", d);
        if (g.has_north) {
            TF_NCCL(g_nccl.Send(p, d * row, ncclInt8, h->rank - 1, comm, s));
            TF_NCCL(g_nccl.Recv(p - d * row, d * row, ncclInt8, h->rank - 1, comm, s));
        }
        if (g.has_south) {
            TF_NCCL(g_nccl.Send(p + (size_t)(g.ny - d) * row, d * row, ncclInt8, h->rank + 1, comm, s));
            TF_NCCL(g_nccl.Recv(p + (size_t)g.ny * row, d * row, ncclInt8, h->rank + 1, comm, s));
        }
    }
    TF_NCCL(g_nccl.GroupEnd());
    return 0;
}

// A few bytes each way between neighbouring strips (per-row tables).
int comm_exchange_bytes(tf_handle h, cudaStream_t s, const void* send_north, void* recv_north,
                        const void* send_south, void* recv_south, size_t bytes) {
    if (!h->nccl_comm) return 0;
    const Geom& g = h->g;
    ncclComm_t comm = (ncclComm_t)h->nccl_comm;
    TF_NCCL(g_nccl.GroupStart());
    if (g.has_north) {
        TF_NCCL(g_nccl.Send(send_north, bytes, ncclInt8, h->rank - 1, comm, s));
        TF_NCCL(g_nccl.Recv(recv_north, bytes, ncclInt8, h->rank - 1, comm, s));
    }
    if (g.has_south) {
        TF_NCCL(g_nccl.Send(send_south, bytes, ncclInt8, h->rank + 1, comm, s));
        TF_NCCL(g_nccl.Recv(recv_south, bytes, ncclInt8, h->rank + 1, comm, s));
    }
    TF_NCCL(g_nccl.GroupEnd());
    return 0;
}

// In-place sum of n uint64 counters over all strips (no-op for one strip).
int comm_allreduce_u64(tf_handle h, cudaStream_t s, unsigned long long* dev, int n) {
    if (!h->nccl_comm) return 0;
    TF_NCCL(g_nccl.AllReduce(dev, dev, (size_t)n, ncclUint64, ncclSum, (ncclComm_t)h->nccl_comm, s));
    return 0;
}

// Every strip's block of rows into one tall raster that all ranks hold:
// rank r's n_rows[r] rows of row_bytes bytes land at full + first_row[r] * row_bytes.
int comm_allgather_rows(tf_handle h, cudaStream_t s, const void* own, void* full, size_t row_bytes,
                        const int* first_row, const int* n_rows) {
    TF_CHECK(h->nccl_comm != nullptr, "comm_allgather_rows: no communicator");
    ncclComm_t comm = (ncclComm_t)h->nccl_comm;
    TF_NCCL(g_nccl.GroupStart());
    for (int r = 0; r < h->nranks; ++r) {
        if (r == h->rank) continue;
        TF_NCCL(g_nccl.Send(own, (size_t)n_rows[h->rank] * row_bytes, ncclInt8, r, comm, s));
        TF_NCCL(g_nccl.Recv((char*)full + (size_t)first_row[r] * row_bytes, (size_t)n_rows[r] * row_bytes,
                            ncclInt8, r, comm, s));
    }
    TF_NCCL(g_nccl.GroupEnd());
    TF_CUDA(cudaMemcpyAsync((char*)full + (size_t)first_row[h->rank] * row_bytes, own,
                            (size_t)n_rows[h->rank] * row_bytes, cudaMemcpyDeviceToDevice, s));
    return 0;
}

int comm_free(tf_handle h) {
    if (h->nccl_comm && g_nccl.CommDestroy) g_nccl.CommDestroy((ncclComm_t)h->nccl_comm);
    h->nccl_comm = nullptr;
    return 0;
}
