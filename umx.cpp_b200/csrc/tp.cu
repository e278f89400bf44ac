// Target-parallel mode (BASELINE config 5): the 4 stem networks are independent until the Wiener stage
// (src/inference.cpp:70-186 touches only model.*[target] and streaming_lstm_data[target]), so rank j of a
// 2- or 4-GPU group runs its share of the targets and ONE exchange step follows: an all-gather of the
// fc3 masks (per segment 4104 x 2584 fp32 = 42.4 MB per target and track) over NVLink.  Every rank then
// computes the shared Wiener statistics redundantly and filters / inverts only its own stems.
// NCCL is plumbing here and is loaded with dlopen on first use (libnccl.so.2: the copy torch has already
// loaded when the caller is a torch process, the system one otherwise); the library itself does not link it.
#include "session.h"

#include <cstring>
#include <dlfcn.h>
#include <mutex>

namespace umx
{
namespace
{
// the few NCCL entry points used, with the ABI of nccl.h (2.x)
struct NcclUniqueId
{
    char internal[128];
};
typedef int (*GetUniqueIdFn)(NcclUniqueId *);
typedef int (*CommInitRankFn)(void **comm, int nranks, NcclUniqueId id, int rank);
typedef int (*AllGatherFn)(const void *send, void *recv, size_t count, int dtype, void *comm, cudaStream_t st);
typedef int (*CommDestroyFn)(void *comm);
typedef const char *(*GetErrorStringFn)(int);
constexpr int kNcclFloat32 = 7; // ncclFloat32 in nccl.h

struct Nccl
{
    void *h = nullptr;
    GetUniqueIdFn get_unique_id = nullptr;
    CommInitRankFn comm_init_rank = nullptr;
    AllGatherFn all_gather = nullptr;
    CommDestroyFn comm_destroy = nullptr;
    GetErrorStringFn err = nullptr;
};

Nccl *nccl()
{
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        n.h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!n.h)
            return;
        n.get_unique_id = (GetUniqueIdFn)dlsym(n.h, "ncclGetUniqueId");
        n.comm_init_rank = (CommInitRankFn)dlsym(n.h, "ncclCommInitRank");
        n.all_gather = (AllGatherFn)dlsym(n.h, "ncclAllGather");
        n.comm_destroy = (CommDestroyFn)dlsym(n.h, "ncclCommDestroy");
        n.err = (GetErrorStringFn)dlsym(n.h, "ncclGetErrorString");
    });
    return (n.h && n.get_unique_id && n.comm_init_rank && n.all_gather && n.comm_destroy) ? &n : nullptr;
}

int check(int rc, const char *what)
{
    if (rc == 0)
        return 0;
    Nccl *n = nccl();
    set_error(std::string(what) + " failed: " + ((n && n->err) ? n->err(rc) : "NCCL error") + " (" + std::to_string(rc) + ")");
    return -1;
}
} // namespace

int tp_unique_id(void *id128)
{
    Nccl *n = nccl();
    UMX_CHECK(n != nullptr, "target-parallel mode needs libnccl.so.2, which could not be loaded");
    return check(n->get_unique_id(static_cast<NcclUniqueId *>(id128)), "ncclGetUniqueId");
}

int tp_init(Session *s, const void *id128, int rank, int world)
{
    UMX_CHECK(world == 1 || world == 2 || world == 4, "target-parallel mode runs on 1, 2 or 4 GPUs (4 targets)");
    UMX_CHECK(rank >= 0 && rank < world, "target-parallel: bad rank");
    tp_destroy(s);
    s->tp_rank = rank;
    s->tp_world = world;
    s->tp_nt = kTargets / world;
    s->tp_t0 = rank * s->tp_nt;
    if (world == 1)
        return 0;
    Nccl *n = nccl();
    UMX_CHECK(n != nullptr, "target-parallel mode needs libnccl.so.2, which could not be loaded");
    UMX_CUDA_OK(cudaSetDevice(s->device));
    NcclUniqueId id;
    std::memcpy(&id, id128, sizeof id);
    return check(n->comm_init_rank(&s->tp_comm, world, id, rank), "ncclCommInitRank");
}

int tp_allgather_masks(Session *s, float *mask, cudaStream_t st)
{
    Nccl *n = nccl();
    UMX_CHECK(n != nullptr && s->tp_comm != nullptr, "target-parallel: communicator is not initialised");
    // in place: this rank's targets are one contiguous slice of the [4][Mcap][4104] mask buffer
    const size_t count = (size_t)s->tp_nt * s->Mcap * kMaskLd;
    s->launches++;
    return check(n->all_gather(mask + (size_t)s->tp_t0 * s->Mcap * kMaskLd, mask, count, kNcclFloat32, s->tp_comm, st),
                 "ncclAllGather");
}

void tp_destroy(Session *s)
{
    if (s->tp_comm)
    {
        if (Nccl *n = nccl())
            n->comm_destroy(s->tp_comm);
        s->tp_comm = nullptr;
    }
    s->tp_rank = 0;
    s->tp_world = 1;
    s->tp_t0 = 0;
    s->tp_nt = kTargets;
}

} // namespace umx
