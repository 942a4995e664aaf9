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
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <mutex>
#include <vector>

struct gsb_window;
struct gsb_comm {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1, device = 0;
    std::vector<gsb_window*> windows;      // live peer windows (orphaned, not freed, if the communicator goes first)
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
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
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
                    bind(g_nccl.AllGather, "ncclAllGather") &&
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


// ---- peer windows: the exchange written as our own kernel over NVLink peer memory ------------------------------------
//
// A window is one cudaMalloc per rank (a 512-byte header of flags + the running means) that every other rank of the node
// maps: cudaIpc handles between processes, raw pointers + cudaDeviceEnablePeerAccess between the devices of one process
// (gsb_comm_init_all).  The all-reduce is then a two-shot kernel: rank r owns the r-th slice of the element range, reads
// that slice from all G windows (G-1 of them through NVLink), adds the G values in rank order and stores the sum into
// all G windows.  Every element is summed by exactly one rank in a fixed order and broadcast as bits, so replicas stay
// bit-identical by construction.  Two system-scope flag barriers bracket it: `ready` (my running means are complete,
// set at kernel entry; the kernel is stream-ordered after the last backward chain) and `done` (I have finished reading
// and writing every window), which the small wait kernel in front of applyGradients polls.  Flags carry the exchange's
// sequence number, so they are never cleared.

constexpr int GSB_MAX_RANKS = 16;
constexpr size_t WIN_HEADER = 512;            // bytes in front of the data: ready[16] | pad | done[16] at +128 | counter at +256
constexpr int WIN_DONE = 32, WIN_CTR = 64;    // u32 indices inside the header
constexpr long long SPIN_LIMIT = 6000000000ll;  // ~3 s of SM clocks: a peer that never arrives ends in an error, not a hang

struct PeerArgs {
    float* data[GSB_MAX_RANKS];
    uint32_t* flags[GSB_MAX_RANKS];
    unsigned long long n4;            // float4 elements in the exchanged range
    int rank, world;
    uint32_t seq;
    uint32_t* host_err;               // mapped pinned word
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void spin_until(const uint32_t* flag, uint32_t seq, uint32_t* host_err) {
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(flag) - seq) < 0) {
        if (clock64() - t0 > SPIN_LIMIT) { *host_err = 1u; break; }
        __nanosleep(64);
    }
}

template <int G>
__global__ void __launch_bounds__(512) peer_allreduce_kernel(PeerArgs a) {
    __shared__ bool s_last;
    const int tid = threadIdx.x;
    const int world = G > 0 ? G : a.world;
    uint32_t* mine = a.flags[a.rank];
    if (blockIdx.x == 0 && tid < world) {
        __threadfence_system();
        st_release_sys(a.flags[tid] + a.rank, a.seq);                 // ready: my running means are complete
    }
    if (tid < world) spin_until(mine + tid, a.seq, a.host_err);
    __syncthreads();
    const unsigned long long per = (a.n4 + world - 1) / world;
    const unsigned long long lo = per * a.rank, hi = min(a.n4, lo + per);
    for (unsigned long long i = lo + (unsigned long long)blockIdx.x * blockDim.x + tid; i < hi; i += (unsigned long long)gridDim.x * blockDim.x) {
        float4 v[G > 0 ? G : GSB_MAX_RANKS];
#pragma unroll
        for (int k = 0; k < (G > 0 ? G : GSB_MAX_RANKS); k++)
            if (k < world) v[k] = __ldcg(reinterpret_cast<const float4*>(a.data[k]) + i);
        float4 acc = v[0];
#pragma unroll
        for (int k = 1; k < (G > 0 ? G : GSB_MAX_RANKS); k++)
            if (k < world) { acc.x += v[k].x; acc.y += v[k].y; acc.z += v[k].z; acc.w += v[k].w; }
#pragma unroll
        for (int k = 0; k < (G > 0 ? G : GSB_MAX_RANKS); k++)
            if (k < world) __stcg(reinterpret_cast<float4*>(a.data[k]) + i, acc);
    }
    // done: every store of every CTA is visible system-wide before the last CTA tells the peers
    __threadfence_system();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(mine + WIN_CTR, 1u) == gridDim.x - 1;
    __syncthreads();
    if (s_last) {
        __threadfence_system();
        if (tid < world) st_release_sys(a.flags[tid] + WIN_DONE + a.rank, a.seq);
        if (tid == 0) mine[WIN_CTR] = 0u;
    }
}

__global__ void peer_wait_kernel(const uint32_t* done, int world, uint32_t seq, uint32_t* host_err) {
    if ((int)threadIdx.x < world) spin_until(done + threadIdx.x, seq, host_err);
}

struct WinMsg {
    long long pid;
    unsigned long long ptr;
    unsigned long long host;
    int device, ok;
    cudaIpcMemHandle_t handle;
};

unsigned long long host_hash() {
    char name[256] = "";
    gethostname(name, sizeof(name) - 1);
    unsigned long long h = 1469598103934665603ull;
    for (const char* c = name; *c; ++c) h = (h ^ (unsigned char)*c) * 1099511628211ull;
    return h;
}

}  // namespace

struct gsb_window {
    gsb_comm* c = nullptr;
    char* base = nullptr;                       // header + data, this rank's allocation
    size_t floats = 0;
    void* peer_base[GSB_MAX_RANKS] = {};
    bool opened[GSB_MAX_RANKS] = {};            // mapped with cudaIpcOpenMemHandle (must be closed)
    bool mapped = false;                        // every rank mapped every peer: the peer kernel may run
    uint32_t seq = 0;
    uint32_t* host_err = nullptr;               // pinned, mapped
    uint32_t* host_err_dev = nullptr;
};

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
    for (gsb_window* w : c->windows) w->c = nullptr;      // their owners destroy them later without the barrier
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

int gsb_comm_window_create(gsb_comm* c, size_t floats, gsb_window** out) {
    if (!c || !out || floats == 0) return gsb_fail_msg("gsb_comm_window_create: invalid argument");
    if (c->world > GSB_MAX_RANKS) return gsb_fail_msg("gsb_comm_window_create: more ranks than a node holds");
    gsb_window* w = new gsb_window();
    w->c = c; w->floats = (floats + 3) & ~(size_t)3;
    GSB_CUDA_OK(cudaMalloc(&w->base, WIN_HEADER + w->floats * sizeof(float)));
    GSB_CUDA_OK(cudaMemset(w->base, 0, WIN_HEADER));
    GSB_CUDA_OK(cudaHostAlloc(&w->host_err, sizeof(uint32_t), cudaHostAllocMapped | cudaHostAllocPortable));
    *w->host_err = 0u;
    GSB_CUDA_OK(cudaHostGetDevicePointer(&w->host_err_dev, w->host_err, 0));
    GSB_CUDA_OK(cudaDeviceSynchronize());
    w->peer_base[c->rank] = w->base;
    c->windows.push_back(w);
    *out = w;
    if (c->world == 1) return GSB_OK;

    // ---- tell every rank where this window lives (collective) ----
    const char* env = getenv("GSB_PEER_EXCHANGE");
    WinMsg mine{};
    mine.pid = (long long)getpid(); mine.ptr = (unsigned long long)w->base; mine.host = host_hash();
    mine.device = c->device;
    mine.ok = !(env && env[0] == '0');
    if (mine.ok && cudaIpcGetMemHandle(&mine.handle, w->base) != cudaSuccess) { cudaGetLastError(); mine.ok = 0; }
    cudaStream_t s;
    GSB_CUDA_OK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    WinMsg* d_msgs = nullptr;
    GSB_CUDA_OK(cudaMalloc(&d_msgs, sizeof(WinMsg) * (size_t)c->world));
    GSB_CUDA_OK(cudaMemcpyAsync(d_msgs + c->rank, &mine, sizeof(WinMsg), cudaMemcpyHostToDevice, s));
    GSB_NCCL_OK(g_nccl.AllGather(d_msgs + c->rank, d_msgs, sizeof(WinMsg), ncclChar, c->comm, s));
    std::vector<WinMsg> msgs((size_t)c->world);
    GSB_CUDA_OK(cudaMemcpyAsync(msgs.data(), d_msgs, sizeof(WinMsg) * (size_t)c->world, cudaMemcpyDeviceToHost, s));
    GSB_CUDA_OK(cudaStreamSynchronize(s));

    bool ok = true;
    for (int k = 0; k < c->world && ok; k++) ok = msgs[(size_t)k].ok && msgs[(size_t)k].host == mine.host;
    for (int k = 0; k < c->world && ok; k++) {
        if (k == c->rank) continue;
        const WinMsg& m = msgs[(size_t)k];
        if (m.pid == mine.pid) {          // another device of this process: its pointer is valid here once peer access is on
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, c->device, m.device) != cudaSuccess || !can) { ok = false; break; }
            const cudaError_t e = cudaDeviceEnablePeerAccess(m.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = false;
            cudaGetLastError();
            w->peer_base[k] = (void*)m.ptr;
        } else if (cudaIpcOpenMemHandle(&w->peer_base[k], m.handle, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess) {
            w->opened[k] = true;
        } else {
            cudaGetLastError();
            ok = false;
        }
    }
    // ---- every rank must take the same path: one failed mapping anywhere sends all of them to ncclAllReduce ----
    float* d_bad = reinterpret_cast<float*>(d_msgs);
    const float bad = ok ? 0.0f : 1.0f;
    GSB_CUDA_OK(cudaMemcpyAsync(d_bad, &bad, sizeof(float), cudaMemcpyHostToDevice, s));
    GSB_NCCL_OK(g_nccl.AllReduce(d_bad, d_bad, 1, ncclFloat, ncclSum, c->comm, s));
    float total = 1.0f;
    GSB_CUDA_OK(cudaMemcpyAsync(&total, d_bad, sizeof(float), cudaMemcpyDeviceToHost, s));
    GSB_CUDA_OK(cudaStreamSynchronize(s));
    w->mapped = total == 0.0f;
    if (!w->mapped)
        for (int k = 0; k < c->world; k++)
            if (w->opened[k]) { cudaIpcCloseMemHandle(w->peer_base[k]); w->opened[k] = false; }
    cudaFree(d_msgs);
    cudaStreamDestroy(s);
    return GSB_OK;
}

int gsb_comm_window_destroy(gsb_window* w) {
    if (!w) return GSB_OK;
    // This device idle means its wait kernel has seen `done` from every peer for the last exchange: no kernel anywhere
    // still touches this rank's window.
    cudaDeviceSynchronize();
    bool ipc = false;
    for (int k = 0; k < GSB_MAX_RANKS; k++)
        if (w->opened[k]) { cudaIpcCloseMemHandle(w->peer_base[k]); ipc = true; }
    int rc = GSB_OK;
    if (ipc && w->c && w->c->comm) {
        // Between processes nobody may free a window a peer still maps: a one-float all-reduce is the barrier between
        // close and free.  (Devices of one process hold raw pointers, nothing to close: their windows may go one by one.)
        float* d = reinterpret_cast<float*>(w->base);
        if (g_nccl.AllReduce(d, d, 1, ncclFloat, ncclSum, w->c->comm, nullptr) != ncclSuccess) rc = GSB_ERR_CUDA;
        cudaDeviceSynchronize();
    }
    if (w->c) w->c->windows.erase(std::remove(w->c->windows.begin(), w->c->windows.end(), w), w->c->windows.end());
    cudaFree(w->base);
    cudaFreeHost(w->host_err);
    delete w;
    return rc;
}

float* gsb_window_data(gsb_window* w) { return w ? reinterpret_cast<float*>(w->base + WIN_HEADER) : nullptr; }
size_t gsb_window_floats(const gsb_window* w) { return w ? w->floats : 0; }
int gsb_window_peer_mapped(const gsb_window* w) { return w && w->mapped ? 1 : 0; }

int gsb_window_allreduce_sum(gsb_window* w, size_t count, void* stream) {
    if (!w || !w->c) return gsb_fail_msg("gsb_window_allreduce_sum: null window");
    if (count > w->floats) return gsb_fail_msg("gsb_window_allreduce_sum: count exceeds the window");
    if (w->c->world == 1 || count == 0) return GSB_OK;
    if (!w->mapped) return gsb_comm_allreduce_sum(w->c, gsb_window_data(w), count, stream);
    if (*w->host_err) { gsb_fail_msg("gsb_window_allreduce_sum: a peer never arrived at an earlier exchange (timed out)"); return GSB_ERR_STATE; }
    PeerArgs a{};
    for (int k = 0; k < w->c->world; k++) {
        a.flags[k] = reinterpret_cast<uint32_t*>(w->peer_base[k]);
        a.data[k] = reinterpret_cast<float*>(static_cast<char*>(w->peer_base[k]) + WIN_HEADER);
    }
    a.n4 = (count + 3) / 4;           // the window is padded to whole float4s; the pad floats are summed harmlessly
    a.rank = w->c->rank; a.world = w->c->world; a.seq = ++w->seq; a.host_err = w->host_err_dev;
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned long long per = (a.n4 + a.world - 1) / a.world;
    const int grid = (int)std::max<unsigned long long>(1, std::min<unsigned long long>(148 * 2, (per + 511) / 512));
    switch (a.world) {
        case 2: peer_allreduce_kernel<2><<<grid, 512, 0, s>>>(a); break;
        case 4: peer_allreduce_kernel<4><<<grid, 512, 0, s>>>(a); break;
        case 8: peer_allreduce_kernel<8><<<grid, 512, 0, s>>>(a); break;
        default: peer_allreduce_kernel<0><<<grid, 512, 0, s>>>(a); break;
    }
    peer_wait_kernel<<<1, 32, 0, s>>>(a.flags[a.rank] + WIN_DONE, a.world, a.seq, a.host_err);
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

}  // extern "C"
