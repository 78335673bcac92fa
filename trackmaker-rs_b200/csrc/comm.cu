// comm.cu — the one collective of the engine: streams shard over the GPUs of a box with no
// data-path traffic; after a batch the 8 decode counters are summed (and a step time maximised) over the ranks
// with NCCL over NVLink / NVSwitch. One process per GPU, one communicator per engine.
//
// libnccl is opened at run time (dlopen "libnccl.so.2", or the path in TMK_NCCL_LIB), so libtmkphy.so loads on a
// box without NCCL and a single-GPU caller never touches it. Only the handful of entry points below are bound;
// their prototypes are NCCL's public C API (nccl.h 2.x: ncclGetUniqueId, ncclCommInitRank, ncclAllReduce,
// ncclCommDestroy, ncclGetErrorString).
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <string>

#include "engine.hpp"

namespace {

constexpr int kNcclUniqueIdBytes = 128;  // NCCL_UNIQUE_ID_BYTES
struct NcclId { char internal[kNcclUniqueIdBytes]; };
typedef void* NcclComm;
// ncclDataType_t / ncclRedOp_t values of nccl.h
constexpr int kNcclUint64 = 5, kNcclFloat64 = 8, kNcclSum = 0, kNcclMax = 2;

struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    std::string err;
};

NcclApi& api() {
    static NcclApi a;
    return a;
}

bool load_nccl() {
    NcclApi& a = api();
    if (a.lib) return true;
    const char* names[] = {getenv("TMK_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        if (!n || !*n) continue;
        a.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (a.lib) break;
        a.err = dlerror() ? dlerror() : "dlopen failed";
    }
    if (!a.lib) return false;
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(a.lib, "ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(a.lib, "ncclCommInitRank"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(a.lib, "ncclAllReduce"));
    a.AllGather = reinterpret_cast<decltype(a.AllGather)>(dlsym(a.lib, "ncclAllGather"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(a.lib, "ncclCommDestroy"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(a.lib, "ncclGetErrorString"));
    if (!a.GetUniqueId || !a.CommInitRank || !a.AllReduce || !a.AllGather || !a.CommDestroy || !a.GetErrorString) {
        a.err = "libnccl lacks a required symbol";
        dlclose(a.lib);
        a.lib = nullptr;
        return false;
    }
    return true;
}

int nccl_fail(tmk_engine* e, const char* what, int rc) {
    return tmk_fail(e, TMK_E_CUDA, std::string(what) + ": " + (api().GetErrorString ? api().GetErrorString(rc) : "nccl error"));
}

}  // namespace

extern "C" {

int tmk_comm_unique_id(void* out128) {
    if (!out128) return TMK_E_INVALID;
    if (!load_nccl()) return TMK_E_CUDA;
    NcclId id;
    if (api().GetUniqueId(&id) != 0) return TMK_E_CUDA;
    std::memcpy(out128, &id, sizeof id);
    return TMK_OK;
}

int tmk_comm_init(tmk_engine* e, const void* unique_id128, int rank, int world) {
    if (!e || !unique_id128 || world < 1 || rank < 0 || rank >= world) return tmk_fail(e, TMK_E_INVALID, "tmk_comm_init: bad argument");
    if (e->comm) return tmk_fail(e, TMK_E_INVALID, "tmk_comm_init: the engine already has a communicator");
    if (!load_nccl()) return tmk_fail(e, TMK_E_CUDA, "tmk_comm_init: cannot load libnccl (" + api().err + ")");
    CK(cudaSetDevice(e->device));
    NcclId id;
    std::memcpy(&id, unique_id128, sizeof id);
    NcclComm comm = nullptr;
    const int rc = api().CommInitRank(&comm, world, id, rank);
    if (rc != 0) return nccl_fail(e, "ncclCommInitRank", rc);
    e->comm = comm;
    e->comm_rank = rank;
    e->comm_world = world;
    CK(e->d_comm.reserve(sizeof(uint64_t) * (64 + 8 + 8 * 64)));
    return TMK_OK;
}

int tmk_comm_size(const tmk_engine* e) { return e && e->comm ? e->comm_world : 1; }

// counters[0..n): summed over the ranks (n <= 32). Without a communicator the call is the identity.
int tmk_comm_allreduce_sum_u64(tmk_engine* e, uint64_t* counters, uint32_t n) {
    if (!e || (!counters && n) || n > 32) return tmk_fail(e, TMK_E_INVALID, "tmk_comm_allreduce_sum_u64: bad argument");
    if (!e->comm || n == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    uint64_t* d = static_cast<uint64_t*>(e->d_comm.p);
    CK(cudaMemcpyAsync(d, counters, n * sizeof(uint64_t), cudaMemcpyHostToDevice, e->stream));
    const int rc = api().AllReduce(d, d, n, kNcclUint64, kNcclSum, e->comm, e->stream);
    if (rc != 0) return nccl_fail(e, "ncclAllReduce", rc);
    CK(cudaMemcpyAsync(counters, d, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

// values[0..n): maximum over the ranks (n <= 32): the device time of a step is the slowest rank's.
int tmk_comm_allreduce_max_f64(tmk_engine* e, double* values, uint32_t n) {
    if (!e || (!values && n) || n > 32) return tmk_fail(e, TMK_E_INVALID, "tmk_comm_allreduce_max_f64: bad argument");
    if (!e->comm || n == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    double* d = reinterpret_cast<double*>(static_cast<uint64_t*>(e->d_comm.p) + 32);
    CK(cudaMemcpyAsync(d, values, n * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    const int rc = api().AllReduce(d, d, n, kNcclFloat64, kNcclMax, e->comm, e->stream);
    if (rc != 0) return nccl_fail(e, "ncclAllReduce", rc);
    CK(cudaMemcpyAsync(values, d, n * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

// all[r * n .. r * n + n) = rank r's mine[0..n) (n <= 8; SURVEY §8e: per-GPU record counts after a batch). Without a
// communicator all = mine.
int tmk_comm_allgather_u64(tmk_engine* e, const uint64_t* mine, uint32_t n, uint64_t* all) {
    if (!e || !mine || !all || n == 0 || n > 8) return tmk_fail(e, TMK_E_INVALID, "tmk_comm_allgather_u64: bad argument");
    if (!e->comm) {
        for (uint32_t i = 0; i < n; ++i) all[i] = mine[i];
        return TMK_OK;
    }
    if (e->comm_world > 64) return tmk_fail(e, TMK_E_INVALID, "tmk_comm_allgather_u64: more than 64 ranks");
    CK(cudaSetDevice(e->device));
    CK(e->d_comm.reserve(sizeof(uint64_t) * (64 + 8 + 8 * 64)));
    uint64_t* src = static_cast<uint64_t*>(e->d_comm.p) + 64;
    uint64_t* dst = src + 8;
    CK(cudaMemcpyAsync(src, mine, n * sizeof(uint64_t), cudaMemcpyHostToDevice, e->stream));
    const int rc = api().AllGather(src, dst, n, kNcclUint64, e->comm, e->stream);
    if (rc != 0) return nccl_fail(e, "ncclAllGather", rc);
    CK(cudaMemcpyAsync(all, dst, size_t(n) * e->comm_world * sizeof(uint64_t), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

// the engine's accumulated decode counters (tmk_stats) summed over all ranks
int tmk_stats_allreduce(tmk_engine* e, uint64_t counters[TMK_STAT_COUNT]) {
    if (!e || !counters) return tmk_fail(e, TMK_E_INVALID, "tmk_stats_allreduce: null argument");
    for (int i = 0; i < TMK_STAT_COUNT; ++i) counters[i] = e->stats[i];
    return tmk_comm_allreduce_sum_u64(e, counters, TMK_STAT_COUNT);
}

// all ranks have reached this point and the engine's stream is idle
int tmk_comm_barrier(tmk_engine* e) {
    if (!e) return TMK_E_INVALID;
    uint64_t one = 1;
    const int rc = tmk_comm_allreduce_sum_u64(e, &one, 1);
    if (rc != TMK_OK) return rc;
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_comm_destroy(tmk_engine* e) {
    if (!e) return TMK_E_INVALID;
    if (e->comm) {
        cudaSetDevice(e->device);
        cudaStreamSynchronize(e->stream);
        api().CommDestroy(e->comm);
        e->comm = nullptr;
        e->comm_world = 1;
        e->comm_rank = 0;
    }
    return TMK_OK;
}

}  // extern "C"
