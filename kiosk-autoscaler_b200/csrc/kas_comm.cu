// kas_comm.cu -- the one collective of the path: box-wide integer totals.
//
// Keys and resources are sharded over the GPUs of a box with no data-path exchange; what is
// exchanged is a handful of integers at the end -- the uint64 counts of tally_queues
// (autoscaler.py:60-77) and the [S][3] pod-tick / patch / final sums of the replayed scale loop
// (autoscaler.py:244-260) -- by ONE ncclAllReduce(sum) over NVLink.  Integer sums are order
// independent, so every world size gives the same bits.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy a host process such as PyTorch
// has already loaded, else the system one), so libkas.so itself links against nothing but the
// CUDA runtime and loads on machines without NCCL; the entry points then return
// KAS_ERR_UNSUPPORTED.  The few NCCL declarations used are restated here (stable C ABI).
#include "kas_common.cuh"

#include <dlfcn.h>
#include <mutex>
#include <string.h>

namespace kas {

struct NcclUniqueId { char internal[KAS_COMM_ID_BYTES]; };
typedef void *NcclComm;
constexpr int kNcclSuccess = 0;
constexpr int kNcclInt64 = 4;    // ncclInt64
constexpr int kNcclSum = 0;      // ncclSum

struct NcclApi {
    int (*GetUniqueId)(NcclUniqueId *);
    int (*CommInitRank)(NcclComm *, int, NcclUniqueId, int);
    int (*AllReduce)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t);
    int (*CommDestroy)(NcclComm);
    const char *(*GetErrorString)(int);
    bool ok;
};

static NcclApi g_nccl = {nullptr, nullptr, nullptr, nullptr, nullptr, false};
static std::once_flag g_nccl_once;

static const NcclApi &nccl() {
    std::call_once(g_nccl_once, [] {
        void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        g_nccl.GetUniqueId = reinterpret_cast<decltype(g_nccl.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
        g_nccl.CommInitRank = reinterpret_cast<decltype(g_nccl.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
        g_nccl.AllReduce = reinterpret_cast<decltype(g_nccl.AllReduce)>(dlsym(h, "ncclAllReduce"));
        g_nccl.CommDestroy = reinterpret_cast<decltype(g_nccl.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
        g_nccl.GetErrorString = reinterpret_cast<decltype(g_nccl.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
        g_nccl.ok = g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.AllReduce && g_nccl.CommDestroy;
    });
    return g_nccl;
}

#define KAS_NCCL_TRY(expr)                                                                  \
    do {                                                                                    \
        const int _r = (expr);                                                              \
        if (_r != kNcclSuccess) {                                                           \
            ::kas::set_text_error(#expr, api.GetErrorString ? api.GetErrorString(_r) : "NCCL error"); \
            return KAS_ERR_CUDA;                                                            \
        }                                                                                   \
    } while (0)

}  // namespace kas

using namespace kas;

extern "C" {

int kas_comm_unique_id(uint8_t *id) {
    if (!id) return KAS_ERR_INVALID;
    const NcclApi &api = nccl();
    if (!api.ok) return KAS_ERR_UNSUPPORTED;
    NcclUniqueId u;
    KAS_NCCL_TRY(api.GetUniqueId(&u));
    memcpy(id, u.internal, KAS_COMM_ID_BYTES);
    return KAS_OK;
}

int kas_comm_init(void **comm, int n_ranks, int rank, const uint8_t *id) {
    if (!comm || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return KAS_ERR_INVALID;
    const NcclApi &api = nccl();
    if (!api.ok) return KAS_ERR_UNSUPPORTED;
    NcclUniqueId u;
    memcpy(u.internal, id, KAS_COMM_ID_BYTES);
    NcclComm c = nullptr;
    KAS_NCCL_TRY(api.CommInitRank(&c, n_ranks, u, rank));
    *comm = c;
    return KAS_OK;
}

int kas_allreduce_i64(void *comm, int64_t *d_buf, size_t n, void *stream) {
    if (!comm || (!d_buf && n)) return KAS_ERR_INVALID;
    if (n == 0) return KAS_OK;
    if (!aligned(d_buf, 8)) return KAS_ERR_ALIGNMENT;
    const NcclApi &api = nccl();
    if (!api.ok) return KAS_ERR_UNSUPPORTED;
    KAS_NCCL_TRY(api.AllReduce(d_buf, d_buf, n, kNcclInt64, kNcclSum, static_cast<NcclComm>(comm),
                               static_cast<cudaStream_t>(stream)));
    return KAS_OK;
}

int kas_comm_destroy(void *comm) {
    if (!comm) return KAS_OK;
    const NcclApi &api = nccl();
    if (!api.ok) return KAS_ERR_UNSUPPORTED;
    KAS_NCCL_TRY(api.CommDestroy(static_cast<NcclComm>(comm)));
    return KAS_OK;
}

}  // extern "C"
