// tmk_internal.cuh — shared definitions of libtmkphy's CUDA translation units (sm_100a only).
//
// Data layout in HBM:
//   samples      float[total]             all captures of a batch, stream s at streams[s].offset
//   blk_info     float4[n_blocks]         per block of TMK_BLK windows: (min E, max E, max corr^2,
//                                          number of candidate windows as float)
//   cand_bits    uint32[n_blocks * 15]    1 bit per window: corr_approx >= thr * CAND_RATIO
//   recs/payload tmk_frame_rec[] / u8[]   verdict records + decoded payload bytes
//
// "Exact" device helpers use __fadd_rn/__fmul_rn/__fdiv_rn/__fsqrt_rn so that nvcc never
// contracts them into FMAs: they reproduce the reference's f32 operation order
// (src/phy/decoder.rs, src/phy/line_coding.rs) bit for bit.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/tmkphy.h"

// cudaFuncSetAttribute is per DEVICE and costs ~25 us: a call site does it once per device (an engine may sit on any of them)
struct TmkPerDevice {
    std::atomic<uint64_t> mask{0};
    bool done(int dev) const { return (mask.load(std::memory_order_acquire) >> (dev & 63)) & 1ull; }
    void set(int dev) { mask.fetch_or(1ull << (dev & 63), std::memory_order_release); }
};


#ifdef TMK_SCAN_ARM_A        // build of the round-1 CUDA-core scan (scan_fast.cu), kept for the A/B of DESIGN.md §3.1
#define TMK_BLK 480          // windows per block (candidate bitmap / energy summary granularity):
                             // 4 lanes x 120 windows of the fast scan kernel, 15 whole bitmap words
#define TMK_BLK_WORDS 15     // TMK_BLK / 32
#define TMK_TILE_BLKS 8      // blocks per scan tile
#else                        // tensor-core scan (scan_tc.cu): rows of 64 windows, 4 rows per block (the four rows one
#define TMK_BLK 256          // epilogue warp holds for one `b`: block statistics need no exchange between warps),
#define TMK_BLK_WORDS 8      // 128 rows per tile
#define TMK_TILE_BLKS 32
#endif
#define TMK_TILE (TMK_BLK * TMK_TILE_BLKS)  // windows per scan tile
#define TMK_MAX_P 1024       // longest supported preamble in samples
#define TMK_MAX_LEVELS 512   // longest supported preamble in levels
#define TMK_CAND_RATIO 0.9444444f  // candidate filter keeps corr_approx >= thr * (0.85 / 0.9)

// Pattern of one (line_code, spl, preamble_bytes) class, resident in device memory.
struct TmkPattern {
    int32_t line_code;
    uint32_t spl;
    uint32_t preamble_bytes;
    uint32_t P;          // preamble samples
    uint32_t L;          // preamble levels (P / spl)
    uint32_t sync_len;   // samples_for_bits(8)
    uint32_t margin;     // samples_for_bits(1)
    uint32_t hdr_samples;  // samples_for_bits(56)
    float pe;            // sqrt(sum p^2)           decoder.rs:46-50
    float sync_energy;   // sqrt(sum sync^2)        decoder.rs:167-171
    float lev_thr;       // smallest level SUM s with fl(s / spl) > 1e-6f (line_coding.rs:268: |avg| > 1e-6 <=> |sum| >= lev_thr)
    float levels[TMK_MAX_LEVELS];  // +1/-1 per level
};

// Per-stream work description for the scan / walk kernels.
struct TmkStreamDev {
    uint64_t offset;      // first sample (multiple of 4)
    uint64_t n_samples;
    uint64_t blk_base;    // first block of this stream in blk_info / cand_bits
    uint32_t tile_base;   // first scan tile of this stream
    uint32_t n_tiles;
    uint32_t pattern;     // index into the pattern table
    int32_t local_addr;
    float threshold;
    uint32_t max_frame_bytes;
    uint64_t chunk;       // 0 = one shot
    // carried PhyDecoder state (links.cu): buffer_offset at entry, and DecoderState::Decoding(frame_start)
    // with the correlation of its detection, or init_fs == TMK_NO_FRAME for DecoderState::Searching
    uint64_t init_q;
    uint64_t init_fs;
    float init_corr;
    uint32_t pad_;
};
#define TMK_NO_FRAME 0xffffffffffffffffull
// PhyDecoder state after the call (decoder.rs:16-33): buffer_offset, Decoding(frame_start) or TMK_NO_FRAME
struct TmkLinkState {
    uint64_t off;
    uint64_t fs;
    float corr;
    uint32_t pad_;
};

struct TmkCounters {
    unsigned long long n_recs;
    unsigned long long n_payload;
    unsigned int next_stream;  // persistent walk warps pull stream indices from here
    unsigned int pad_;
    unsigned long long stats[TMK_STAT_COUNT];
    unsigned long long dbg[8];  // walk diagnostics (tmk_debug_counters)
};

// ---- cluster pass (walk.cu): per-candidate results computed in parallel before the per-stream walk
struct TmkCandEntry {  // one candidate window of a block, in window order (32 bytes: two 16-byte loads)
    uint16_t wl;       // window offset inside the block
    uint16_t status;   // verdict of the speculative record (TMK_FRAME_*), or TMK_ENT_NOSPEC
    float D;           // exact AVX-order dot product   (decoder.rs:381-417)
    float Ec;          // accurately summed window energy
    uint32_t spec;     // speculative frame record of a detection at this window, or TMK_ENT_INVALID
    uint32_t consumed; // samples the frame consumes from the window (what the walk needs to go on without the record)
    int32_t fs_off;    // frame_start - window
    uint32_t len;      // payload length of the header (chunked feeding: how many samples decode_frame waited for)
    uint32_t pad_;
};
#define TMK_ENT_INVALID 0xffffffffu
#define TMK_ENT_NOSPEC 0xffffu
struct TmkFast {
    const uint32_t* cblk_list;       // global ids of the blocks with candidates (appended by the scan kernels)
    unsigned int* cblk_count;
    uint2* blk_res;                  // per global block with candidates: x = (first entry << 6) | count, or TMK_ENT_INVALID;
                                     // y = gain (float bits): max_w (window energies of the block up to w) / S_w
    TmkCandEntry* entries;
    uint32_t entry_cap;
    tmk_frame_rec* spec;             // speculative verdict records (payload already in the payload buffer)
    uint32_t spec_cap;
    unsigned int* pool;              // [0] entries used, [1] spec records used
    unsigned int* item_next;         // work counter of the cluster pass (one per chunk)
};

// ---- segmented walk (walk.cu): a long one-shot capture is walked by several warps at once. Walker j starts a search
// at the first window of segment j as if the decoder were Searching there; the state of the reference's decoder at a
// search entry is its position alone (one-shot feeding: every sample is buffered, the rolling energy is re-seeded at the
// entry), so from the first search entry that a walker shares with the walk coming from the segments before it, its
// records ARE that walk's. Pass 1: every walker runs to its first search entry at or behind the end of its segment.
// Pass 2: it goes on from there until its search entry is one of the first entries of the walker that owns the position
// (or the capture ends). Resolve: from walker 0 (exact by definition) along those meeting points; the records of a
// walker before its meeting point and of walkers that were skipped are dropped.
struct TmkSegItem {
    uint32_t stream;      // the capture
    uint32_t seg;         // segment index inside it
    uint32_t first_item;  // item of the capture's segment 0
    uint32_t n_seg;       // segments of the capture
    uint64_t q0, q1;      // first search entry; pause at the first search entry >= q1 (the last segment: never)
    uint64_t seg_len;     // windows per segment
};
#define TMK_SEG_HEADS 8
struct TmkSegState {
    uint64_t q;           // search entry at which the walker paused (pass 1) / met another walker (pass 2)
    uint32_t ord;         // records emitted so far
    uint32_t state;       // 0 ended (capture exhausted), 1 paused, 2 met walker sync_item at its entry sync_ord
    uint32_t sync_item, sync_ord;
    uint32_t n_head, pad_;
    uint64_t head[TMK_SEG_HEADS];  // search entry after record 0, 1, ... of pass 1
};
struct TmkSeg {
    const TmkSegItem* items;  // nullptr: one walker per capture (the plain walk)
    TmkSegState* st;
    uint32_t pass;
    uint32_t pad_;
};

__host__ __device__ inline uint64_t tmk_sfb(int line_code, uint32_t spl, uint64_t nbits) {
    // LineCode::samples_for_bits  line_coding.rs:125-127 (Manchester), :295-300 (4B5B)
    return line_code == TMK_LINE_MANCHESTER ? nbits * spl * 2 : ((nbits + 3) / 4) * 5 * spl;
}

#ifdef __CUDACC__
// ---------------------------------------------------------------- exact f32 helpers
__device__ __forceinline__ float ex_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float ex_sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ex_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float ex_div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float ex_sqrt(float a) { return __fsqrt_rn(a); }

// rolling window energy update with clamp, decoder.rs:224-233
__device__ __forceinline__ float ex_roll(float e, float leaving, float entering) {
    e = ex_add(ex_sub(e, ex_mul(leaving, leaving)), ex_mul(entering, entering));
    return e < 0.0f ? 0.0f : e;
}

// correlation of one window given its dot product and window energy, decoder.rs:142-147
__device__ __forceinline__ float ex_corr(float dot, float energy, float pe) {
    if (energy < 1e-6f) return 0.0f;
    return ex_div(dot, ex_mul(ex_sqrt(energy), pe));
}

__device__ __forceinline__ float max_nan(float a, float b) {
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

// CRC-8 poly 0x07 init 0 (crc.rs:7-22), one byte step
__device__ __forceinline__ uint32_t crc8_step(uint32_t crc, uint32_t byte) {
    crc ^= byte;
#pragma unroll
    for (int k = 0; k < 8; ++k) crc = (crc & 0x80u) ? ((crc << 1) ^ 0x07u) & 0xFFu : (crc << 1) & 0xFFu;
    return crc;
}
#endif

// kernel launchers (each returns the number of kernels it launched)
int tmk_launch_tile_dir(const TmkStreamDev* streams, uint32_t n_streams, uint32_t n_tiles, uint32_t* tile_stream,
                        cudaStream_t st);
// kinds: bit 0 = 4B5B/spl 3/2 bytes streams present, bit 1 = Manchester/3/2, bit 2 = any other shape
int tmk_launch_scan(const float* samples, uint64_t total_samples, const TmkStreamDev* streams, uint32_t n_streams,
                    const uint32_t* tile_stream, uint32_t tile0, uint32_t n_tiles, const TmkPattern* patterns,
                    uint32_t kinds, float4* blk_info, uint32_t* cand_bits, uint32_t* cblk_list, unsigned int* cblk_count,
                    cudaStream_t st, void* tile_desc, float* dbg_dot = nullptr, bool build_dir = false);
// build_dir: the descriptor kernel also fills tile_stream (tile0 == 0, all tiles): tmk_launch_tile_dir is not needed then
// cblk_list / cblk_count (nullable): global ids of the blocks with candidates; tile_desc: device scratch of
// tmk_scan_tile_desc_bytes(n_tiles) bytes; dbg_dot (nullable): the filter's correlation numerators per window
size_t tmk_scan_tile_desc_bytes(uint32_t n_tiles);
int tmk_launch_scan_generic(const float* samples, const TmkStreamDev* streams, const uint32_t* tile_stream,
                            uint32_t tile0, uint32_t n_tiles, const TmkPattern* patterns, float4* blk_info,
                            uint32_t* cand_bits, uint32_t* cblk_list, unsigned int* cblk_count, cudaStream_t st);
// walks streams [s0, s1); `next` is a zeroed work counter owned by this launch
// cluster pass over the blocks listed in f.cblk_list (one-shot streams; others are marked invalid)
int tmk_launch_cluster(const float* samples, const TmkStreamDev* streams, const uint32_t* tile_stream,
                       const TmkPattern* patterns, const float4* blk_info, const uint32_t* cand_bits, TmkFast f,
                       uint8_t* payload, uint64_t payload_cap, TmkCounters* counters, bool beside_walk, cudaStream_t st);
int tmk_launch_walk(const float* samples, const TmkStreamDev* streams, uint32_t s0, uint32_t s1, unsigned int* next,
                    TmkFast fast,
                    const TmkPattern* patterns, const float4* blk_info, const uint32_t* cand_bits,
                    tmk_frame_rec* recs, uint4* rec_ord, uint64_t rec_cap, uint8_t* payload,
                    uint64_t payload_cap, uint32_t* stream_cnt, TmkCounters* counters, TmkLinkState* link_out,
                    cudaStream_t st, TmkSeg seg = TmkSeg{nullptr, nullptr, 0, 0});
// segmented walk: valid record range per walker (first_valid / n_valid), verdict tallies of the valid records, and the
// ordering of the valid records into (capture, time) order
int tmk_launch_seg_resolve(const TmkSegItem* items, const TmkSegState* st, uint32_t n_items, uint32_t* first_valid,
                           uint32_t* n_valid, const tmk_frame_rec* recs, const uint4* rec_ord, uint64_t n_slots, uint64_t rec_cap,
                           const tmk_frame_rec* spec, TmkCounters* counters, cudaStream_t stream);
int tmk_launch_seg_chain(const TmkSegItem* items, const TmkSegState* st, uint32_t n_items, uint32_t* first_valid, uint32_t* n_valid,
                         cudaStream_t stream);
int tmk_launch_seg_order(const tmk_frame_rec* recs_in, const uint4* rec_ord, uint64_t n_slots, const uint64_t* base,
                         const tmk_frame_rec* spec, const TmkSegItem* items, const uint32_t* first_valid,
                         tmk_frame_rec* recs_out, cudaStream_t stream);
int tmk_launch_replay_probe(const float* samples, const TmkStreamDev* streams, const TmkPattern* patterns, uint64_t seed,
                            const uint64_t* ws, uint32_t n, float* out, unsigned long long* counts, cudaStream_t st);
int tmk_launch_order_small(const uint32_t* stream_cnt, uint64_t* stream_base, uint32_t n_streams, const tmk_frame_rec* recs_in,
                           const uint4* rec_ord, const TmkCounters* counters, uint64_t slot_cap, const tmk_frame_rec* spec,
                           tmk_frame_rec* recs_out, uint64_t first_cap, tmk_frame_rec* recs_out_rest, uint64_t out_cap, uint64_t* n_exact,
                           cudaStream_t st);
int tmk_launch_order(const tmk_frame_rec* recs_in, const uint4* rec_ord, uint64_t n_recs,
                     const uint32_t* stream_cnt, uint64_t* stream_base, uint32_t n_streams,
                     const tmk_frame_rec* spec, tmk_frame_rec* recs_out, cudaStream_t st);
int tmk_launch_excl_scan(const uint32_t* cnt, uint64_t* base, uint32_t n, cudaStream_t st);
int tmk_walk_init();
int tmk_walk_max_frame_bytes();
int tmk_launch_correlate(const float* samples, const TmkStreamDev* streams, uint32_t n_streams,
                         const TmkPattern* patterns, float* corr_out, cudaStream_t st);
int tmk_launch_busy(const float* samples, const TmkStreamDev* streams, uint32_t n_streams, int8_t* busy,
                    cudaStream_t st);
// 16-bit PCM ingest / dump (misc.cu): codec.rs:121-125 and dump.rs:37-40
int tmk_launch_pcm16_to_f32(const int16_t* pcm, float* out, uint64_t n, cudaStream_t st);
int tmk_launch_pcm16_encode(const float* samples, int16_t* out, uint64_t n, cudaStream_t st);
int tmk_launch_modulate(const TmkPattern* pattern, const tmk_frame_desc* frames, const uint64_t* frame_off,
                        uint32_t n_frames, const uint8_t* payload, float* out, cudaStream_t st);
// synthetic captures (synth.cu)
struct TmkSynthImp {      // impairments of tmk_synth_batch_ex (defaults: no multipath, no drift, gaps U[48, 4800))
    uint32_t multipath;
    float drift_ppm;
    uint32_t gap_min, gap_span;  // inter-frame gap = gap_min + U[0, gap_span)
    float amp_min;               // > 0: per-stream amplitude scale U[amp_min, 1]
};
int tmk_launch_synth_noise(uint64_t seed, uint64_t total, float sigma, uint64_t n_per_stream, TmkSynthImp imp,
                           float* samples, cudaStream_t st);
int tmk_launch_synth_plan(const TmkPattern* pattern, int32_t local_addr, uint64_t seed, uint32_t n_streams,
                          uint64_t n_per_stream, TmkSynthImp imp, uint32_t* counts, const uint64_t* base,
                          tmk_frame_rec* frames, uint64_t frame_cap, cudaStream_t st);
int tmk_launch_synth_frames(const TmkPattern* pattern, uint64_t seed, uint64_t n_per_stream, float amplitude,
                            TmkSynthImp imp, tmk_frame_rec* frames, uint64_t n_frames, float* samples, cudaStream_t st);
int tmk_launch_ask_synth(const float* tab, uint64_t seed, uint32_t n_streams, uint64_t n_per_stream, float* samples,
                         unsigned long long* n_frames, cudaStream_t st);
