// gsb_comm_*: the exchange step of multi-GPU training as a native entry point.
//
// The reference is single-GPU; its only cross-view interaction is the linear sum of accumulateGradients
// (src/Trainer.cu:47-77, called :416-419) followed by applyGradients (:428-431).  Sharding the truth cameras
// over G devices therefore needs exactly one sum of the 24 running-mean floats per splat per iteration, between
// accumulate and apply: one ncclAllReduce on the trainer's stream (NVLink 5 / NVSwitch, NVLS when NCCL picks it).
// NCCL delivers bit-identical sums to every rank and apply / densify are deterministic, so replicas stay identical.
//
// NCCL is bound at run time (dlopen of libnccl.so.2, re-using a copy the process already loaded, e.g. the one torch
// ships) so that libgsb200.so itself carries no NCCL dependency: single-GPU callers never touch it.

#include "common.cuh"
#include "../../include/gsb.h"

#include <dlfcn.h>
#include <nccl.h>

#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

struct gsb_comm {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1, device = 0;
};

namespace {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi g_nccl;
std::once_flag g_nccl_once;
char g_nccl_err[256] = "";

template <typename F>
bool bind(F& fn, const char* name) {
    fn = reinterpret_cast<F>(dlsym(g_nccl.handle, name));
    if (!fn) snprintf(g_nccl_err, sizeof(g_nccl_err), "libnccl lacks %s", name);
    return fn != nullptr;
}

void load_nccl() {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names)   // a copy already mapped into the process (torch's) wins: two NCCLs must not coexist
        if ((g_nccl.handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL))) break;
    if (!g_nccl.handle)
        for (const char* n : names)
            if ((g_nccl.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!g_nccl.handle) {
        snprintf(g_nccl_err, sizeof(g_nccl_err), "cannot load libnccl.so.2: %s", dlerror());
        return;
    }
    const bool ok = bind(g_nccl.GetUniqueId, "ncclGetUniqueId") && bind(g_nccl.CommInitRank, "ncclCommInitRank") &&
                    bind(g_nccl.CommInitAll, "ncclCommInitAll") && bind(g_nccl.CommDestroy, "ncclCommDestroy") &&
                    bind(g_nccl.AllReduce, "ncclAllReduce") && bind(g_nccl.Broadcast, "ncclBroadcast") &&
                    bind(g_nccl.GroupStart, "ncclGroupStart") && bind(g_nccl.GroupEnd, "ncclGroupEnd") &&
                    bind(g_nccl.GetVersion, "ncclGetVersion") && bind(g_nccl.GetErrorString, "ncclGetErrorString");
    if (!ok) { dlclose(g_nccl.handle); g_nccl.handle = nullptr; }
}

int need_nccl() {
    std::call_once(g_nccl_once, load_nccl);
    if (!g_nccl.handle) { gsb_fail_msg(g_nccl_err); return GSB_ERR_STATE; }
    return GSB_OK;
}

int nccl_fail(ncclResult_t r, const char* what) {
    char msg[320];
    snprintf(msg, sizeof(msg), "NCCL error %d (%s) in %s", (int)r, g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?", what);
    gsb_fail_msg(msg);
    return GSB_ERR_CUDA;
}

#define GSB_NCCL_OK(expr)                                       \
    do {                                                        \
        ncclResult_t _r = (expr);                               \
        if (_r != ncclSuccess) return nccl_fail(_r, #expr);     \
    } while (0)

}  // namespace

extern "C" {

int gsb_comm_nccl_version(int* version) {
    if (!version) return gsb_fail_msg("gsb_comm_nccl_version: null argument");
    if (int e = need_nccl()) return e;
    GSB_NCCL_OK(g_nccl.GetVersion(version));
    return GSB_OK;
}

int gsb_comm_get_unique_id(void* id_out) {
    static_assert(sizeof(ncclUniqueId) == GSB_COMM_UNIQUE_ID_BYTES, "ncclUniqueId size");
    if (!id_out) return gsb_fail_msg("gsb_comm_get_unique_id: null argument");
    if (int e = need_nccl()) return e;
    ncclUniqueId id;
    GSB_NCCL_OK(g_nccl.GetUniqueId(&id));
    memcpy(id_out, &id, sizeof(id));
    return GSB_OK;
}

int gsb_comm_init(gsb_comm** out, const void* nccl_unique_id, int rank, int world) {
    if (!out || !nccl_unique_id || world <= 0 || rank < 0 || rank >= world) return gsb_fail_msg("gsb_comm_init: invalid argument");
    if (int e = need_nccl()) return e;
    ncclUniqueId id;
    memcpy(&id, nccl_unique_id, sizeof(id));
    gsb_comm* c = new gsb_comm();
    c->rank = rank; c->world = world;
    GSB_CUDA_OK(cudaGetDevice(&c->device));
    ncclResult_t r = g_nccl.CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) { delete c; return nccl_fail(r, "ncclCommInitRank"); }
    *out = c;
    return GSB_OK;
}

int gsb_comm_init_all(gsb_comm** out, int ndev, const int* devices) {
    if (!out || ndev <= 0) return gsb_fail_msg("gsb_comm_init_all: invalid argument");
    if (int e = need_nccl()) return e;
    int have = 0;
    GSB_CUDA_OK(cudaGetDeviceCount(&have));
    if (ndev > have) return gsb_fail_msg("gsb_comm_init_all: more devices requested than visible");
    std::vector<ncclComm_t> comms((size_t)ndev);
    std::vector<int> devs((size_t)ndev);
    for (int i = 0; i < ndev; i++) devs[(size_t)i] = devices ? devices[i] : i;
    GSB_NCCL_OK(g_nccl.CommInitAll(comms.data(), ndev, devs.data()));
    for (int i = 0; i < ndev; i++) {
        gsb_comm* c = new gsb_comm();
        c->comm = comms[(size_t)i]; c->rank = i; c->world = ndev; c->device = devs[(size_t)i];
        out[i] = c;
    }
    return GSB_OK;
}

int gsb_comm_destroy(gsb_comm* c) {
    if (!c) return GSB_OK;
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    delete c;
    return GSB_OK;
}

int gsb_comm_rank(const gsb_comm* c) { return c ? c->rank : 0; }
int gsb_comm_world(const gsb_comm* c) { return c ? c->world : 1; }
int gsb_comm_device(const gsb_comm* c) { return c ? c->device : 0; }

int gsb_comm_allreduce_sum(gsb_comm* c, float* dev, size_t count, void* stream) {
    if (!c || !c->comm) return gsb_fail_msg("gsb_comm_allreduce_sum: null communicator");
    if (count == 0 || c->world == 1) return GSB_OK;
    GSB_NCCL_OK(g_nccl.AllReduce(dev, dev, count, ncclFloat, ncclSum, c->comm, (cudaStream_t)stream));
    return GSB_OK;
}

int gsb_comm_broadcast(gsb_comm* c, void* dev, size_t bytes, int root, void* stream) {
    if (!c || !c->comm) return gsb_fail_msg("gsb_comm_broadcast: null communicator");
    if (bytes == 0 || c->world == 1) return GSB_OK;
    GSB_NCCL_OK(g_nccl.Broadcast(dev, dev, bytes, ncclChar, root, c->comm, (cudaStream_t)stream));
    return GSB_OK;
}

}  // extern "C"
