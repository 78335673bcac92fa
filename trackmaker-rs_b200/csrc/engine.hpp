// engine.hpp — the engine object behind the C ABI and the small host helpers shared by the
// translation units that implement entry points (engine.cu, ask.cu).
#pragma once
#include <atomic>
#include <string>
#include <vector>

#include "tmk_internal.cuh"

struct DevBuf {  // grow-only device scratch
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) {
            e = cudaMalloc(&p, bytes);
            want = bytes;
        }
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct tmk_engine {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;
    cudaStream_t walk_streams[4] = {};  // cluster pass + walk of chunk g run on walk_streams[g % 4] (high priority)
    cudaEvent_t wev[4] = {};
    std::string err;
    std::vector<TmkPattern> patterns;  // host mirror of the device pattern table
    DevBuf d_tile_dir, d_patterns, d_streams, d_blk_info, d_cand, d_recs_tmp, d_rec_ord, d_recs, d_payload, d_cnt, d_base,
        d_counters, d_samples, d_misc, d_misc2, d_next, d_cblk, d_blk_ptr, d_entries, d_spec, d_pool, d_link, d_pcm, d_frag;
    DevBuf d_ask_tab, d_ask_bits, d_ask_l2, d_ask_recs, d_ask_recs2, d_ask_chk, d_ask_tc, d_ask_seg;  // chirp+BPSK path (ask.cu)
    DevBuf d_mcm_tab, d_mcm_recs, d_mcm_off;  // multi-carrier modem (mcm.cu): tables, device records, stream offsets
    bool mcm_ready = false;
    float mcm_demod_ms = 0.0f;
    DevBuf d_seg;  // segmented walk: items, walker states, first valid record per walker
    DevBuf d_small;             // small batches: counters + exit states + ordered records + payload in one block (engine.cu)
    void* h_small = nullptr;    // its pinned host mirror
    size_t h_small_cap = 0;
    void* h_streams = nullptr;  // pinned staging of the stream descriptors (a copy out of pageable memory is staged by the runtime: ~8 us for 256 links)
    size_t h_streams_cap = 0;
    DevBuf d_tile_desc;  // per-tile descriptors of the tensor-core scan (scan_tc.cu)
    DevBuf d_comm, d_wav, d_wavdesc;  // staging of the counter all-reduce (comm.cu); raw WAV bytes + descriptors (wav.cu)
    float* scan_dbg_dot = nullptr;    // set around tmk_scan_probe by tmk_scan_dot_probe (diagnostic)
    void* comm = nullptr;             // ncclComm_t (comm.cu), one per engine
    int comm_rank = 0, comm_world = 1;
    bool ask_ready = false;
    float ask_hscale = 0.0f;  // max |H| / 200 of the chirp spectrum (error bound of the FFT filter)
    cudaEvent_t ev[6] = {};
    cudaEvent_t tev[2] = {};
    cudaEvent_t cev[6] = {};  // copy/compute hand-over of the chunked staging
    float stage_ms[4] = {0, 0, 0, 0};
    bool ev5_valid = false;      // ev[5] (end of the cluster pass) was recorded by the last decode
    bool stage_pending = false;  // small batches leave the event arithmetic to whoever asks (tmk_flush_stage_ms)
    float ask_ms[2] = {0, 0};  // last tmk_ask_decode_batch: scan, walk
    uint64_t launches = 0;
    uint64_t stats[TMK_STAT_COUNT] = {};
    uint64_t dbg[8] = {};
    // set by links.cu around one tmk_decode_batch call: per-stream PhyDecoder state at entry / at exit (host arrays)
    const TmkLinkState* link_init = nullptr;
    TmkLinkState* link_final = nullptr;
};

inline int tmk_fail(tmk_engine* e, int code, const std::string& msg) {
    if (e) e->err = msg;
    return code;
}

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t _err = (call);                                                                 \
        if (_err != cudaSuccess)                                                                   \
            return tmk_fail(e, _err == cudaErrorMemoryAllocation ? TMK_E_NOMEM : TMK_E_CUDA,       \
                            std::string(#call) + ": " + cudaGetErrorString(_err));                 \
    } while (0)

inline bool tmk_is_device_ptr(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// engine.cu: stage times of the last decode out of its events, if nobody has asked yet — before e->ev[] are recorded again
void tmk_flush_stage_ms(tmk_engine* e);
// ask.cu, shared with mcm.cu (the multi-carrier frames use the chirp preamble and the frame envelope of examples/ask.rs)
int tmk_ask_ensure_tables(tmk_engine* e);
int tmk_ask_sync_device(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams, uint32_t n_streams,
                        const float** dx_out, uint64_t* n_recs, uint64_t* dev_cap_out, bool want_bits = true);
// ask_tc.cu: the chirp filter on the tensor cores (marks candidate words in wflag)
int tmk_ask_tc_prepare(tmk_engine* e, const uint64_t* s_off, const uint64_t* s_n, const uint64_t* s_word, uint32_t n_streams,
                       uint32_t* tile_base_out);
int tmk_ask_tc_filter(tmk_engine* e, const float* dx, uint64_t total_samples, uint32_t n_streams, uint32_t tile0, uint32_t tile1,
                      const float* chirp, const float* pchk, uint32_t* wflag, cudaStream_t st);
