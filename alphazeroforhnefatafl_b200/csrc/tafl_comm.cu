// tafl_comm.cu — the one collective of the path: gather of the fixed-stride replay
// records of every rank to the replay-buffer rank over NCCL (NVLink / NVSwitch inside a box).
//
// Games shard by id range, one process per GPU; stepping and searching exchange nothing.  Per move
// each rank holds n_local records of TAFL_REC_STRIDE bytes (tafl_pack_records); ranks may hold
// different numbers of rows, so the counts are all-gathered first and the payload then moves with
// grouped ncclSend / ncclRecv straight into its final offset on the root (no padding, no staging).
//
// libnccl is resolved at RUN time (dlopen), so the library has no link-time dependency on it: a host
// process that already carries NCCL (PyTorch does) shares that copy, a Rust host loads libnccl.so.2
// from the loader path.  Only the handful of entry points used here are bound, by their public
// prototypes (nccl.h); ncclComm_t / ncclUniqueId cross the C ABI as opaque pointers / 128 bytes.
#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "tafl_engine.hpp"

using namespace tafl;

namespace {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;    // ncclSuccess == 0
typedef int ncclDataType_t;  // ncclUint8 == 1, ncclUint32 == 3
constexpr ncclDataType_t kUint8 = 1, kUint32 = 3;

struct NcclApi {
    void* h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;
};

NcclApi& nccl() {
    static NcclApi api;
    if (api.h || !api.err.empty()) return api;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {   // a copy already in the process (PyTorch's) wins
        api.h = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (api.h) break;
    }
    std::string why;   // dlerror() of every failed attempt (the call clears the message, so read it once each)
    auto try_open = [&](const char* path) {
        dlerror();
        void* h = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
        if (!h) {
            const char* de = dlerror();
            why += std::string(why.empty() ? "" : "; ") + path + ": " + (de ? de : "dlopen failed");
        }
        return h;
    };
    if (!api.h) {
        if (const char* p = std::getenv("TAFL_NCCL_LIB")) api.h = try_open(p);
        for (const char* n : names) {
            if (api.h) break;
            api.h = try_open(n);
        }
    }
    if (!api.h) {
        api.err = "libnccl.so.2 not found (set TAFL_NCCL_LIB): " + why;
        return api;
    }
#define TAFL_BIND(field, sym)                                                   \
    do {                                                                        \
        *(void**)(&api.field) = dlsym(api.h, sym);                              \
        if (!api.field) { api.err = std::string("missing NCCL symbol ") + sym; api.h = nullptr; return api; } \
    } while (0)
    TAFL_BIND(GetUniqueId, "ncclGetUniqueId");
    TAFL_BIND(CommInitRank, "ncclCommInitRank");
    TAFL_BIND(CommDestroy, "ncclCommDestroy");
    TAFL_BIND(AllGather, "ncclAllGather");
    TAFL_BIND(Send, "ncclSend");
    TAFL_BIND(Recv, "ncclRecv");
    TAFL_BIND(GroupStart, "ncclGroupStart");
    TAFL_BIND(GroupEnd, "ncclGroupEnd");
    TAFL_BIND(GetErrorString, "ncclGetErrorString");
#undef TAFL_BIND
    return api;
}

int nccl_fail(tafl_engine* e, ncclResult_t r, const char* what) {
    if (e) e->err = std::string(what) + ": " + (nccl().GetErrorString ? nccl().GetErrorString(r) : "NCCL error");
    return TAFL_ERR_CUDA;
}
#define TAFL_NCCL_TRY(e, call)                                   \
    do {                                                         \
        ncclResult_t _r = (call);                                \
        if (_r != 0) return nccl_fail((e), _r, #call);           \
    } while (0)

}  // namespace

extern "C" {

int tafl_nccl_unique_id(void* out128) {
    if (!out128) return TAFL_ERR_INVALID_ARG;
    NcclApi& n = nccl();
    if (!n.h) return TAFL_ERR_CUDA;
    ncclUniqueId id;
    if (n.GetUniqueId(&id) != 0) return TAFL_ERR_CUDA;
    std::memcpy(out128, &id, sizeof(id));
    return TAFL_OK;
}

int tafl_nccl_comm_create(tafl_engine* e, const void* id128, int n_ranks, int rank, void** out_comm) {
    if (!e || !id128 || !out_comm || n_ranks < 1 || rank < 0 || rank >= n_ranks) return TAFL_ERR_INVALID_ARG;
    NcclApi& n = nccl();
    if (!n.h) { e->err = n.err; return TAFL_ERR_CUDA; }
    cudaSetDevice(e->device);
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    ncclComm_t comm = nullptr;
    TAFL_NCCL_TRY(e, n.CommInitRank(&comm, n_ranks, id, rank));
    *out_comm = (void*)comm;
    return TAFL_OK;
}

int tafl_nccl_comm_destroy(void* comm) {
    NcclApi& n = nccl();
    if (!n.h || !comm) return TAFL_ERR_INVALID_ARG;
    return n.CommDestroy((ncclComm_t)comm) == 0 ? TAFL_OK : TAFL_ERR_CUDA;
}

int tafl_gather_records(tafl_engine* e, void* comm_, int n_ranks, int rank, int root, const void* d_send, uint32_t n_local,
                        void* d_recv, uint64_t recv_capacity_rows, uint32_t* h_counts) {
    return tafl_gather_records_on(e, comm_, n_ranks, rank, root, d_send, n_local, d_recv, recv_capacity_rows, h_counts, nullptr);
}

int tafl_gather_records_on(tafl_engine* e, void* comm_, int n_ranks, int rank, int root, const void* d_send, uint32_t n_local,
                           void* d_recv, uint64_t recv_capacity_rows, uint32_t* h_counts, void* stream_) {
    if (!e || !comm_ || n_ranks < 1 || rank < 0 || rank >= n_ranks || root < 0 || root >= n_ranks) return TAFL_ERR_INVALID_ARG;
    if (!h_counts && rank == root) return TAFL_ERR_INVALID_ARG;   // the root needs the counts to post its receives
    if (n_local && !d_send) return TAFL_ERR_INVALID_ARG;
    NcclApi& n = nccl();
    if (!n.h) { e->err = n.err; return TAFL_ERR_CUDA; }
    ncclComm_t comm = (ncclComm_t)comm_;
    cudaSetDevice(e->device);
    const cudaStream_t stream = stream_ ? (cudaStream_t)stream_ : e->stream;
    // 1. row counts of every rank (device staging: slot 0 = own count, slot 1 = all counts)
    uint32_t* d_cnt;
    void* p;
    int rc = scratch(e, 5, (size_t)(n_ranks + 1) * sizeof(uint32_t), &p);
    if (rc) return rc;
    d_cnt = (uint32_t*)p;
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(d_cnt, &n_local, sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
    TAFL_NCCL_TRY(e, n.AllGather(d_cnt, d_cnt + 1, 1, kUint32, comm, stream));
    if (!h_counts) {   // a sending rank that does not ask for the counts: nothing to wait for, the call stays asynchronous
        if (n_local) TAFL_NCCL_TRY(e, n.Send(d_send, (size_t)n_local * TAFL_REC_STRIDE, kUint8, root, comm, stream));
        return TAFL_OK;
    }
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_counts, d_cnt + 1, (size_t)n_ranks * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
    // The root waits here for the slowest rank's count.  A blocking-sync event yields the core instead of spinning on
    // it: eight ranks with their NCCL proxy threads share the host's cores, and a spinning waiter takes one from a
    // rank that still has launches to enqueue.
    if (!e->ev_block) TAFL_CUDA_TRY(e, cudaEventCreateWithFlags(&e->ev_block, cudaEventBlockingSync | cudaEventDisableTiming));
    TAFL_CUDA_TRY(e, cudaEventRecord(e->ev_block, stream));
    TAFL_CUDA_TRY(e, cudaEventSynchronize(e->ev_block));
    uint64_t total = 0;
    for (int r = 0; r < n_ranks; ++r) total += h_counts[r];
    void* d_sink = nullptr;   // root buffer too small: the exchange must still happen (the senders are committed to it)
    if (rank == root && total > 0 && (!d_recv || total > recv_capacity_rows)) {
        TAFL_CUDA_TRY(e, cudaMalloc(&d_sink, (size_t)total * TAFL_REC_STRIDE));
        d_recv = d_sink;
    }
    // 2. payload: every rank -> root, rows land at their final offset (rank order == game-id order).
    // Nothing returns from inside the group: the first error is kept, ncclGroupEnd always runs (an open
    // group would poison every later collective of this thread), then the sink is released.
    cudaError_t cerr = cudaSuccess;
    ncclResult_t nerr = 0;
    const char* nwhat = "";
    uint64_t own_off = 0;
    if (rank == root && n_local) {   // the root's own rows: a plain device copy, outside the group
        for (int r = 0; r < root; ++r) own_off += (uint64_t)h_counts[r] * TAFL_REC_STRIDE;
        cerr = cudaMemcpyAsync((char*)d_recv + own_off, d_send, (size_t)n_local * TAFL_REC_STRIDE, cudaMemcpyDeviceToDevice, stream);
    }
    nerr = n.GroupStart();
    if (nerr) nwhat = "ncclGroupStart";
    else {
        if (rank == root) {
            uint64_t off = 0;
            for (int r = 0; r < n_ranks; ++r) {
                const size_t bytes = (size_t)h_counts[r] * TAFL_REC_STRIDE;
                if (bytes && r != root && !nerr) {
                    nerr = n.Recv((char*)d_recv + off, bytes, kUint8, r, comm, stream);
                    if (nerr) nwhat = "ncclRecv";
                }
                off += bytes;
            }
        } else if (n_local) {
            nerr = n.Send(d_send, (size_t)n_local * TAFL_REC_STRIDE, kUint8, root, comm, stream);
            if (nerr) nwhat = "ncclSend";
        }
        const ncclResult_t gerr = n.GroupEnd();
        if (!nerr && gerr) { nerr = gerr; nwhat = "ncclGroupEnd"; }
    }
    if (d_sink) {
        cudaStreamSynchronize(stream);
        cudaFree(d_sink);
    }
    if (cerr != cudaSuccess) return set_cuda_error(e, cerr, "cudaMemcpyAsync(root rows)");
    if (nerr) return nccl_fail(e, nerr, nwhat);
    if (d_sink) {
        e->err = "tafl_gather_records: receive buffer too small on the root (h_counts holds the rows per rank)";
        return TAFL_ERR_CAPACITY;
    }
    return TAFL_OK;
}

}  // extern "C"
