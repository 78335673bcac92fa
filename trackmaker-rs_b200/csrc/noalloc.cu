// noalloc.cu — tmk_channel_busy_noalloc: mac::is_channel_busy (src/mac/mod.rs:18-27) over a batch of streams for
// callers that may not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc). No counterpart in the reference.
//
// tmk_channel_busy uploads its stream descriptors into grow-only device scratch of the engine (cudaMalloc on first
// use or growth). This variant keeps ALL of its scratch — the descriptors, one "saw a loud sample" flag per stream
// and a staging row for host-side results — in ONE file-scope `static __device__` array of fixed size, which every
// GPU that loads the library holds until the process exits (TMK_NOALLOC_MAX_STREAMS streams, 84 KB). It calls no
// cudaMalloc / cudaFree / cudaMallocHost and no allocating library routine.
//
// The array is shared by every call on a device and keeps the last call's contents: the flags, which must start at
// zero, are cleared at the start of every call; calls that overlap on one GPU (two engines, two host threads) would
// race on it, so the entry point serialises its callers with a process-wide mutex until its stream has drained —
// callers in DIFFERENT processes on one GPU each have their own copy of the array.
//
// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol, not a device address: the address
// the kernels receive comes from cudaGetSymbolAddress, resolved once per device with that device current.
#include <mutex>
#include <string>

#include "engine.hpp"

namespace {

constexpr uint32_t kMaxStreams = TMK_NOALLOC_MAX_STREAMS;
constexpr int kMaxDevices = 64;
constexpr uint32_t kSlice = 65536;  // samples per CTA of the partial kernel

struct BusyScratch {
    uint64_t offset[kMaxStreams];
    uint64_t n_samples[kMaxStreams];
    uint32_t any[kMaxStreams];  // must be zero when the partial kernel starts
    int8_t out[kMaxStreams];    // results of a call whose `busy` is a host pointer
};

static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(BusyScratch)];

// grid (slices of the longest stream, streams): a CTA looks at one slice of one stream
__global__ void __launch_bounds__(256) busy_partial_kernel(const float* __restrict__ samples, BusyScratch* __restrict__ sc) {
    const uint32_t s = blockIdx.y;
    const uint64_t n = sc->n_samples[s];
    const uint64_t lo = uint64_t(blockIdx.x) * kSlice;
    if (n < 20 || lo >= n) return;
    const uint64_t hi = lo + kSlice < n ? lo + kSlice : n;
    const float* __restrict__ x = samples + sc->offset[s];
    int any = 0;
    for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) any |= fabsf(x[i]) > 0.5f;
    any = __syncthreads_or(any);
    if (any && threadIdx.x == 0) atomicOr(&sc->any[s], 1u);
}

__global__ void busy_final_kernel(BusyScratch* __restrict__ sc, uint32_t n_streams, int8_t* __restrict__ busy) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    busy[s] = sc->n_samples[s] < 20 ? int8_t(-1) : int8_t(sc->any[s] ? 1 : 0);
}

std::mutex g_mu;                          // serialises the users of the array (and of the host staging rows below)
BusyScratch* g_addr[kMaxDevices] = {};    // device address of g_noalloc_scratch per device, resolved on first use
uint64_t g_stage[2 * kMaxStreams];        // host staging of the descriptors (static: no heap allocation per call)

}  // namespace

extern "C" int tmk_channel_busy_noalloc(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                                        uint32_t n_streams, int8_t* busy) {
    if (!e || (!samples && total_samples) || !streams || !busy)
        return tmk_fail(e, TMK_E_INVALID, "tmk_channel_busy_noalloc: null argument");
    if (n_streams > kMaxStreams)
        return tmk_fail(e, TMK_E_CAPACITY, "tmk_channel_busy_noalloc: " + std::to_string(n_streams) + " streams exceed the static scratch (TMK_NOALLOC_MAX_STREAMS = " +
                                               std::to_string(kMaxStreams) + "); split the batch or call tmk_channel_busy");
    if (e->device < 0 || e->device >= kMaxDevices) return tmk_fail(e, TMK_E_INVALID, "tmk_channel_busy_noalloc: device index out of range");
    CK(cudaSetDevice(e->device));
    if (total_samples && !tmk_is_device_ptr(samples))
        return tmk_fail(e, TMK_E_INVALID, "tmk_channel_busy_noalloc: samples must be device-resident (staging a host buffer would allocate)");
    if (n_streams == 0) return TMK_OK;

    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_addr[e->device]) {  // once per device, with that device current
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_addr[e->device] = static_cast<BusyScratch*>(p);
    }
    BusyScratch* sc = g_addr[e->device];
    uint64_t longest = 0;
    for (uint32_t s = 0; s < n_streams; ++s) {
        if (streams[s].offset > total_samples || streams[s].n_samples > total_samples - streams[s].offset)
            return tmk_fail(e, TMK_E_INVALID, "tmk_channel_busy_noalloc: stream exceeds the sample buffer");
        g_stage[s] = streams[s].offset;
        g_stage[n_streams + s] = streams[s].n_samples;
        if (streams[s].n_samples > longest) longest = streams[s].n_samples;
    }
    const uint64_t slices = (longest + kSlice - 1) / kSlice;
    if (slices > 0x7fffffffull) return tmk_fail(e, TMK_E_INVALID, "tmk_channel_busy_noalloc: stream too long");
    CK(cudaMemcpyAsync(sc->offset, g_stage, sizeof(uint64_t) * n_streams, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(sc->n_samples, g_stage + n_streams, sizeof(uint64_t) * n_streams, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemsetAsync(sc->any, 0, sizeof(uint32_t) * n_streams, e->stream));  // the previous call's flags are still there
    const bool dev_out = tmk_is_device_ptr(busy);
    int8_t* d_out = dev_out ? busy : sc->out;
    if (slices) {
        busy_partial_kernel<<<dim3(uint32_t(slices), n_streams), 256, 0, e->stream>>>(samples, sc);
        e->launches += 1;
    }
    busy_final_kernel<<<(n_streams + 255) / 256, 256, 0, e->stream>>>(sc, n_streams, d_out);
    e->launches += 1;
    CK(cudaGetLastError());
    if (!dev_out) CK(cudaMemcpyAsync(busy, d_out, n_streams, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));  // the array is free again when the lock is released
    return TMK_OK;
}
