// engine.cu — the C ABI of libtmkphy (include/tmkphy.h): host-side batch plumbing around the
// scan / walk kernels. No CPU implementation of any PHY stage lives here: every entry point that
// computes launches CUDA kernels, and fails with TMK_E_CUDA when no sm_100 device is usable.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "engine.hpp"

namespace {

int fail(tmk_engine* e, int code, const std::string& msg) { return tmk_fail(e, code, msg); }
bool is_device_ptr(const void* p) { return tmk_is_device_ptr(p); }

bool valid_cfg(const tmk_phy_config& c) {
    if (c.line_code != TMK_LINE_MANCHESTER && c.line_code != TMK_LINE_4B5B) return false;
    if (c.spl == 0 || c.preamble_bytes == 0 || c.spl > 64 || c.preamble_bytes > 64) return false;
    const uint64_t P = tmk_sfb(c.line_code, c.spl, 8ull * c.preamble_bytes);
    if (P > TMK_MAX_P || P / c.spl > TMK_MAX_LEVELS) return false;
    if (!(c.threshold > 0.0f) || !(c.threshold <= 1.0f)) return false;
    if (c.local_addr < -1 || c.local_addr > 255) return false;
    if (c.max_frame_bytes > uint32_t(tmk_walk_max_frame_bytes())) return false;
    return true;
}

// LineCode::generate_preamble (line_coding.rs:13-21) as level signs, for either line code
void build_pattern(const tmk_phy_config& c, TmkPattern* p) {
    std::memset(p, 0, sizeof *p);
    p->line_code = c.line_code;
    p->spl = c.spl;
    p->preamble_bytes = c.preamble_bytes;
    std::vector<int> bits;
    static const int k33[8] = {0, 0, 1, 1, 0, 0, 1, 1}, k5a[8] = {0, 1, 0, 1, 1, 0, 1, 0};
    for (uint32_t i = 0; i + 1 < c.preamble_bytes; ++i) bits.insert(bits.end(), k33, k33 + 8);
    bits.insert(bits.end(), k5a, k5a + 8);
    uint32_t L = 0;
    if (c.line_code == TMK_LINE_MANCHESTER) {
        for (int b : bits) {
            p->levels[L++] = b ? -1.0f : 1.0f;
            p->levels[L++] = b ? 1.0f : -1.0f;
        }
    } else {
        static const int enc[16] = {0x1e, 0x09, 0x14, 0x15, 0x0a, 0x0b, 0x0e, 0x0f,
                                    0x12, 0x13, 0x16, 0x17, 0x1a, 0x1b, 0x1c, 0x1d};
        float level = 1.0f;
        for (size_t i = 0; i < bits.size(); i += 4) {
            const int nib = bits[i] << 3 | bits[i + 1] << 2 | bits[i + 2] << 1 | bits[i + 3];
            for (int j = 4; j >= 0; --j) {
                if ((enc[nib] >> j) & 1) level = -level;
                p->levels[L++] = level;
            }
        }
    }
    p->L = L;
    p->P = L * c.spl;
    p->sync_len = uint32_t(tmk_sfb(c.line_code, c.spl, 8));
    p->margin = uint32_t(tmk_sfb(c.line_code, c.spl, 1));
    p->hdr_samples = uint32_t(tmk_sfb(c.line_code, c.spl, 8 * TMK_PHY_HEADER_BYTES));
    // sqrt of a sequential f32 sum of +-1 squares: exact integers, so sqrtf(P) (decoder.rs:46-50,167-171)
    p->pe = std::sqrt(float(p->P));
    p->sync_energy = std::sqrt(float(p->sync_len));
    {   // the f32 quotient sum / spl is monotone in the sum: find where it crosses the 1e-6 gate of the NRZI decoder
        const float dv = float(c.spl);
        volatile float t = 1e-6f * dv;
        while (float(t / dv) > 1e-6f) t = std::nextafterf(t, 0.0f);
        while (!(float(t / dv) > 1e-6f)) t = std::nextafterf(t, 1.0f);
        p->lev_thr = t;
    }
}

// index of the pattern of `c` in the engine's table (uploads the table when it grows)
int pattern_index(tmk_engine* e, const tmk_phy_config& c, uint32_t* out) {
    for (size_t i = 0; i < e->patterns.size(); ++i) {
        const TmkPattern& p = e->patterns[i];
        if (p.line_code == c.line_code && p.spl == c.spl && p.preamble_bytes == c.preamble_bytes) {
            *out = uint32_t(i);
            return TMK_OK;
        }
    }
    TmkPattern p;
    build_pattern(c, &p);
    e->patterns.push_back(p);
    CK(cudaStreamSynchronize(e->stream));
    CK(e->d_patterns.reserve(sizeof(TmkPattern) * std::max<size_t>(16, e->patterns.size())));
    CK(cudaMemcpyAsync(e->d_patterns.p, e->patterns.data(), sizeof(TmkPattern) * e->patterns.size(),
                       cudaMemcpyHostToDevice, e->stream));
    *out = uint32_t(e->patterns.size() - 1);
    return TMK_OK;
}

struct Batch {
    std::vector<TmkStreamDev> sd;
    uint64_t n_tiles = 0, n_blocks = 0;
    uint32_t kinds = 0;  // which scan kernels the batch needs
};

int prepare_batch(tmk_engine* e, uint64_t total_samples, const tmk_stream* streams, uint32_t n_streams, Batch* b) {
    b->sd.resize(n_streams);
    uint64_t tiles = 0;
    for (uint32_t s = 0; s < n_streams; ++s) {
        const tmk_stream& st = streams[s];
        if (!valid_cfg(st.cfg)) return fail(e, TMK_E_INVALID, "stream " + std::to_string(s) + ": invalid tmk_phy_config");
        if (st.offset % 4 != 0) return fail(e, TMK_E_INVALID, "stream offset must be a multiple of 4 samples");
        if (st.offset > total_samples || st.n_samples > total_samples - st.offset)
            return fail(e, TMK_E_INVALID, "stream " + std::to_string(s) + " exceeds the sample buffer");
        TmkStreamDev& d = b->sd[s];
        uint32_t pi = 0;
        int rc = pattern_index(e, st.cfg, &pi);
        if (rc != TMK_OK) return rc;
        const uint64_t P = e->patterns[pi].P;
        const uint64_t nwin = st.n_samples >= P ? st.n_samples - P + 1 : 0;
        d.offset = st.offset;
        d.n_samples = st.n_samples;
        d.n_tiles = uint32_t((nwin + TMK_TILE - 1) / TMK_TILE);
        if (tiles + d.n_tiles > 0x7fffffffull) return fail(e, TMK_E_INVALID, "batch too large (tile count)");
        d.tile_base = uint32_t(tiles);
        d.blk_base = tiles * TMK_TILE_BLKS;
        d.pattern = pi;
        d.local_addr = st.cfg.local_addr;
        d.threshold = st.cfg.threshold;
        d.max_frame_bytes = st.cfg.max_frame_bytes;
        d.chunk = st.cfg.chunk;
        d.init_q = 0;
        d.init_fs = TMK_NO_FRAME;
        d.init_corr = 0.f;
        d.pad_ = 0;
        tiles += d.n_tiles;
        const TmkPattern& pp = e->patterns[pi];
        if (d.n_tiles) b->kinds |= (pp.spl == 3 && pp.preamble_bytes == 2) ? (pp.line_code == TMK_LINE_4B5B ? 1u : 2u) : 4u;
    }
    b->n_tiles = tiles;
    b->n_blocks = tiles * TMK_TILE_BLKS;
    return TMK_OK;
}

// device view of caller memory: the pointer itself when it is device memory, else a staged copy
int stage_in(tmk_engine* e, DevBuf& buf, const void* src, size_t bytes, const void** out) {
    if (bytes == 0) {
        *out = nullptr;
        return TMK_OK;
    }
    if (is_device_ptr(src)) {
        *out = src;
        return TMK_OK;
    }
    CK(buf.reserve(bytes));
    CK(cudaMemcpyAsync(buf.p, src, bytes, cudaMemcpyHostToDevice, e->stream));
    *out = buf.p;
    return TMK_OK;
}

}  // namespace

void tmk_flush_stage_ms(tmk_engine* e) {
    if (!e || !e->stage_pending) return;
    e->stage_pending = false;
    float cms = 0.f;
    if (e->ev5_valid && cudaEventElapsedTime(&cms, e->ev[2], e->ev[5]) == cudaSuccess) e->dbg[4] = uint64_t(cms * 1000.f);
    cudaEventElapsedTime(&e->stage_ms[3], e->ev[0], e->ev[1]);
    cudaEventElapsedTime(&e->stage_ms[0], e->ev[1], e->ev[2]);
    cudaEventElapsedTime(&e->stage_ms[1], e->ev[2], e->ev[3]);
    cudaEventElapsedTime(&e->stage_ms[2], e->ev[3], e->ev[4]);
    cudaGetLastError();
}

extern "C" {

int tmk_abi_version(void) { return TMK_ABI_VERSION; }

int tmk_engine_create(int device, tmk_engine** out) {
    if (!out) return TMK_E_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
        cudaGetLastError();
        return TMK_E_CUDA;  // no CPU fallback: without a CUDA device there is no engine
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major != 10) return TMK_E_CUDA;  // sm_100a only
    if (cudaSetDevice(device) != cudaSuccess) return TMK_E_CUDA;
    tmk_engine* e = new tmk_engine();
    e->device = device;
    bool ok = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking) == cudaSuccess;
    {
        int lo_pri = 0, hi_pri = 0;  // numerically lowest = highest priority
        cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri);
        for (int i = 0; ok && i < 4; ++i)
            ok = cudaStreamCreateWithPriority(&e->walk_streams[i], cudaStreamNonBlocking, hi_pri) == cudaSuccess &&
                 cudaEventCreateWithFlags(&e->wev[i], cudaEventDisableTiming) == cudaSuccess;
    }
    for (int i = 0; ok && i < 6; ++i) ok = cudaEventCreate(&e->ev[i]) == cudaSuccess;
    for (int i = 0; ok && i < 2; ++i) ok = cudaEventCreate(&e->tev[i]) == cudaSuccess;
    for (int i = 0; ok && i < 6; ++i) ok = cudaEventCreateWithFlags(&e->cev[i], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && tmk_walk_init() == 0;
    ok = ok && e->d_counters.reserve(sizeof(TmkCounters)) == cudaSuccess;
    if (!ok) {
        tmk_engine_destroy(e);
        return TMK_E_CUDA;
    }
    *out = e;
    return TMK_OK;
}

void tmk_engine_destroy(tmk_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    tmk_comm_destroy(e);
    DevBuf* bufs[] = {&e->d_comm, &e->d_wav, &e->d_wavdesc, &e->d_tile_desc, &e->d_tile_dir, &e->d_patterns, &e->d_streams, &e->d_blk_info, &e->d_cand, &e->d_recs_tmp, &e->d_rec_ord,
                      &e->d_recs, &e->d_payload, &e->d_cnt, &e->d_base, &e->d_counters, &e->d_samples,
                      &e->d_misc, &e->d_misc2, &e->d_next, &e->d_cblk, &e->d_blk_ptr, &e->d_entries, &e->d_spec, &e->d_pool, &e->d_link, &e->d_pcm, &e->d_frag,
                      &e->d_ask_tab, &e->d_ask_bits, &e->d_ask_l2,
                      &e->d_ask_recs, &e->d_ask_recs2, &e->d_ask_chk, &e->d_ask_tc, &e->d_ask_seg, &e->d_seg, &e->d_mcm_tab, &e->d_mcm_recs, &e->d_mcm_off};
    for (DevBuf* b : bufs) b->release();
    e->d_small.release();
    if (e->h_small) cudaFreeHost(e->h_small);
    e->h_small = nullptr;
    if (e->h_streams) cudaFreeHost(e->h_streams);
    e->h_streams = nullptr;
    for (int i = 0; i < 6; ++i)
        if (e->ev[i]) cudaEventDestroy(e->ev[i]);
    for (int i = 0; i < 2; ++i)
        if (e->tev[i]) cudaEventDestroy(e->tev[i]);
    for (int i = 0; i < 6; ++i)
        if (e->cev[i]) cudaEventDestroy(e->cev[i]);
    if (e->stream) cudaStreamDestroy(e->stream);
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    for (int i = 0; i < 4; ++i) {
        if (e->walk_streams[i]) cudaStreamDestroy(e->walk_streams[i]);
        if (e->wev[i]) cudaEventDestroy(e->wev[i]);
    }
    delete e;
}

const char* tmk_last_error(const tmk_engine* e) { return e ? e->err.c_str() : "null engine"; }

void tmk_default_config(int32_t line_code, int32_t local_addr, tmk_phy_config* out) {
    if (!out) return;
    out->line_code = line_code;
    out->spl = TMK_DEFAULT_SPL;
    out->preamble_bytes = TMK_DEFAULT_PREAMBLE_BYTES;
    out->local_addr = local_addr;
    out->threshold = TMK_DEFAULT_THRESHOLD;
    out->max_frame_bytes = TMK_DEFAULT_MAX_FRAME_BYTES;
    out->chunk = 0;
}

uint64_t tmk_samples_for_bits(const tmk_phy_config* cfg, uint64_t n_bits) {
    return cfg ? tmk_sfb(cfg->line_code, cfg->spl, n_bits) : 0;
}

uint64_t tmk_preamble_len(const tmk_phy_config* cfg) {
    return cfg ? tmk_sfb(cfg->line_code, cfg->spl, 8ull * cfg->preamble_bytes) : 0;
}

uint64_t tmk_modulated_len(const tmk_phy_config* cfg, const tmk_frame_desc* frames, uint32_t n_frames,
                           uint32_t gap_samples) {
    if (!cfg || (!frames && n_frames)) return 0;
    uint64_t total = 0;
    for (uint32_t i = 0; i < n_frames; ++i) {
        total += tmk_preamble_len(cfg) + tmk_sfb(cfg->line_code, cfg->spl, 8ull * (TMK_PHY_HEADER_BYTES + frames[i].len));
        if (i + 1 < n_frames) total += gap_samples;  // between frames only, encoder.rs:81-83
    }
    return total;
}

int tmk_modulate(tmk_engine* e, const tmk_phy_config* cfg, const tmk_frame_desc* frames, uint32_t n_frames,
                 const uint8_t* payload, uint64_t payload_len, uint32_t gap_samples, float* out_samples,
                 uint64_t out_cap, uint64_t* out_len) {
    if (!e || !cfg || (!frames && n_frames) || !out_len) return fail(e, TMK_E_INVALID, "tmk_modulate: null argument");
    if (!valid_cfg(*cfg)) return fail(e, TMK_E_INVALID, "tmk_modulate: invalid config");
    CK(cudaSetDevice(e->device));
    std::vector<uint64_t> off(n_frames + 1, 0);
    const uint64_t P = tmk_preamble_len(cfg);
    for (uint32_t i = 0; i < n_frames; ++i) {
        if (frames[i].len > 65535u || frames[i].payload_off > payload_len || frames[i].len > payload_len - frames[i].payload_off)
            return fail(e, TMK_E_INVALID, "tmk_modulate: frame " + std::to_string(i) + " payload out of range");
        uint64_t flen = P + tmk_sfb(cfg->line_code, cfg->spl, 8ull * (TMK_PHY_HEADER_BYTES + frames[i].len));
        off[i + 1] = off[i] + flen + (i + 1 < n_frames ? gap_samples : 0);
    }
    const uint64_t total = off[n_frames];
    *out_len = total;
    if (total > out_cap) return fail(e, TMK_E_CAPACITY, "tmk_modulate: output buffer too small");
    if (total == 0) return TMK_OK;
    if (!out_samples) return fail(e, TMK_E_INVALID, "tmk_modulate: null output");
    uint32_t pi = 0;
    int rc = pattern_index(e, *cfg, &pi);
    if (rc != TMK_OK) return rc;
    // frame table + offsets + payload -> device
    const size_t fbytes = sizeof(tmk_frame_desc) * n_frames, obytes = sizeof(uint64_t) * (n_frames + 1);
    CK(e->d_misc.reserve(fbytes + obytes + 16));
    char* dm = static_cast<char*>(e->d_misc.p);
    CK(cudaMemcpyAsync(dm, off.data(), obytes, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(dm + obytes, frames, fbytes, cudaMemcpyHostToDevice, e->stream));
    const void* d_payload = nullptr;
    rc = stage_in(e, e->d_payload, payload, payload_len, &d_payload);
    if (rc != TMK_OK) return rc;
    float* d_out = out_samples;
    const bool dev_out = is_device_ptr(out_samples);
    if (!dev_out) {
        CK(e->d_samples.reserve(total * sizeof(float)));
        d_out = static_cast<float*>(e->d_samples.p);
    }
    e->launches += tmk_launch_modulate(static_cast<const TmkPattern*>(e->d_patterns.p) + pi,
                                       reinterpret_cast<const tmk_frame_desc*>(dm + obytes),
                                       reinterpret_cast<const uint64_t*>(dm), n_frames,
                                       static_cast<const uint8_t*>(d_payload), d_out, e->stream);
    CK(cudaGetLastError());
    if (!dev_out) CK(cudaMemcpyAsync(out_samples, d_out, total * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

// tmk_decode_batch / tmk_decode_batch_pcm16: `samples_v` is f32, or int16 PCM that is widened on the device
// (codec.rs:121-125) chunk by chunk in front of the scan.
static int decode_batch_impl(tmk_engine* e, const void* samples_v, bool pcm16, uint64_t total_samples, const tmk_stream* streams,
                             uint32_t n_streams, tmk_frame_rec* recs, uint64_t rec_cap, uint64_t* n_recs, uint8_t* payload,
                             uint64_t payload_cap, uint64_t* payload_len) {
    const float* samples = pcm16 ? nullptr : static_cast<const float*>(samples_v);
    const int16_t* pcm = pcm16 ? static_cast<const int16_t*>(samples_v) : nullptr;
    if (!e || (!samples_v && total_samples) || (!streams && n_streams) || !n_recs || !payload_len ||
        (!recs && rec_cap) || (!payload && payload_cap))
        return fail(e, TMK_E_INVALID, "tmk_decode_batch: null argument");
    *n_recs = 0;
    *payload_len = 0;
    if (n_streams == 0) return TMK_OK;
    static const bool trace = getenv("TMK_TRACE_HOST") != nullptr;  // host-side time line of a small batch on stderr (us)
    const auto t_in = std::chrono::steady_clock::now();
    double tr[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // trace stamps: [0] batch prepared, [1] buffers reserved, [2] scan, [3] cluster, [4] walk enqueued
    auto us_since = [&](std::chrono::steady_clock::time_point t) {
        return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t).count();
    };
    CK(cudaSetDevice(e->device));
    e->ev5_valid = false;
    e->stage_pending = false;  // (this call records the events again: stage times nobody asked for are dropped, not computed)
    Batch b;
    int rc = prepare_batch(e, total_samples, streams, n_streams, &b);
    if (rc != TMK_OK) return rc;
    if (trace) tr[0] = us_since(t_in);
    // carried PhyDecoder state of the links module (links.cu): entry state in, exit state out
    const TmkLinkState* link_init = e->link_init;
    TmkLinkState* link_final = e->link_final;
    e->link_init = nullptr;
    e->link_final = nullptr;
    TmkLinkState* d_link = nullptr;
    if (link_init)
        for (uint32_t s = 0; s < n_streams; ++s) {
            b.sd[s].init_q = link_init[s].off;
            b.sd[s].init_fs = link_init[s].fs;
            b.sd[s].init_corr = link_init[s].corr;
        }
    if (link_final) {
        CK(e->d_link.reserve(sizeof(TmkLinkState) * n_streams));
        d_link = static_cast<TmkLinkState*>(e->d_link.p);
    }

    CK(e->d_streams.reserve(sizeof(TmkStreamDev) * n_streams));
    {
        // (every call ends with a synchronisation of e->stream, so the staging buffer is free again here)
        const size_t sb = sizeof(TmkStreamDev) * n_streams;
        const void* src = b.sd.data();
        if (sb <= (8u << 20)) {
            if (e->h_streams_cap < sb) {
                if (e->h_streams) cudaFreeHost(e->h_streams);
                e->h_streams = nullptr;
                e->h_streams_cap = 0;
                if (cudaHostAlloc(&e->h_streams, sb + sb / 2 + 4096, cudaHostAllocDefault) == cudaSuccess) e->h_streams_cap = sb + sb / 2 + 4096;
                else cudaGetLastError();
            }
            if (e->h_streams_cap >= sb) {
                std::memcpy(e->h_streams, b.sd.data(), sb);
                src = e->h_streams;
            }
        }
        CK(cudaMemcpyAsync(e->d_streams.p, src, sb, cudaMemcpyHostToDevice, e->stream));
    }
    CK(cudaEventRecord(e->ev[0], e->stream));
    // Host-resident captures are staged in chunks of whole streams: the copy of chunk g+1 (copy stream)
    // overlaps the scan + walk of chunk g (compute stream). Device-resident captures are one chunk.
    const bool host_samples = total_samples && !is_device_ptr(samples_v);
    struct Chunk { uint32_t s0, s1, t0, t1; uint64_t lo, hi; };
    std::vector<Chunk> chunks;
    {
        bool ordered = true;  // streams in increasing, non-overlapping offset order
        for (uint32_t s = 1; s < n_streams && ordered; ++s)
            ordered = streams[s].offset >= streams[s - 1].offset + streams[s - 1].n_samples;
        // samples per chunk: 1 GB. The walks of the chunks share one stream and each is a ~5-8 ms serial chain, so
        // chunks must be large enough for a walk to hide behind the next copy (measured: 8 M -> 1.8, 48 M -> 9.6,
        // 256 M -> 13.0, one chunk -> 12.2 Gsamples/s end to end at C2)
        uint64_t target = 256ull << 20;
        // (device-resident captures: one chunk unless TMK_RESIDENT_CHUNK_MSAMPLES asks otherwise — measured at C2,
        //  the walk CTAs fill the register file, so chunk walks and later scans take turns instead of overlapping:
        //  one chunk 14.05 ms, 256 M 13.9, 128 M 17.5, 64 M 23.4 ms per step)
        const char* ev = getenv(host_samples ? "TMK_CHUNK_MSAMPLES" : "TMK_RESIDENT_CHUNK_MSAMPLES");
        if (ev) target = std::max<uint64_t>(1, strtoull(ev, nullptr, 10)) << 20;
        if (!ordered || total_samples <= target || (!host_samples && !ev)) {
            chunks.push_back({0, n_streams, 0, uint32_t(b.n_tiles), 0, total_samples});
        } else {
            uint32_t s0 = 0;
            while (s0 < n_streams) {
                uint32_t s1 = s0;
                const uint64_t lo = streams[s0].offset;
                uint64_t hi = lo;
                while (s1 < n_streams && (s1 == s0 || streams[s1].offset + streams[s1].n_samples - lo <= target)) {
                    hi = streams[s1].offset + streams[s1].n_samples;
                    ++s1;
                }
                const uint32_t t0 = b.sd[s0].tile_base;
                const uint32_t t1 = s1 < n_streams ? b.sd[s1].tile_base : uint32_t(b.n_tiles);
                chunks.push_back({s0, s1, t0, t1, lo, hi});
                s0 = s1;
            }
        }
    }
    const float* dx = samples;
    const int16_t* dpcm = pcm;
    if (host_samples || pcm16) {
        CK(e->d_samples.reserve(total_samples * sizeof(float)));
        dx = static_cast<const float*>(e->d_samples.p);
    }
    if (pcm16) {
        if (reinterpret_cast<uintptr_t>(pcm) % 8 != 0) return fail(e, TMK_E_INVALID, "pcm buffer must be 8-byte aligned");
        if (host_samples) {
            CK(e->d_pcm.reserve(total_samples * sizeof(int16_t)));
            dpcm = static_cast<const int16_t*>(e->d_pcm.p);
        }
    }
    CK(e->d_blk_info.reserve(sizeof(float4) * std::max<uint64_t>(1, b.n_blocks)));
    CK(e->d_cand.reserve(sizeof(uint32_t) * TMK_BLK_WORDS * std::max<uint64_t>(1, b.n_blocks)));
    // Segmented walk (tmk_internal.cuh): when the batch has too few captures to fill the GPU with one warp per capture, long
    // one-shot captures are cut into segments walked at once (BASELINE config 4: an hour-long capture is 172.8 M samples)
    std::vector<TmkSegItem> items;
    std::vector<uint32_t> item_base(n_streams + 1, 0);
    bool seg_mode = false;
    {
        uint64_t target_walkers = 148ull * 24, seg_min = 131072;
        if (const char* ev = getenv("TMK_WALK_SEGMENTS")) target_walkers = strtoull(ev, nullptr, 10);  // 0 / 1: off
        if (const char* ev = getenv("TMK_WALK_SEG_MIN")) seg_min = std::max<uint64_t>(1024, strtoull(ev, nullptr, 10));
        if (!link_init && !link_final && total_samples && uint64_t(n_streams) * 2 <= target_walkers) {
            items.reserve(size_t(target_walkers) + n_streams);
            for (uint32_t s = 0; s < n_streams; ++s) {
                item_base[s] = uint32_t(items.size());
                const TmkStreamDev& d = b.sd[s];
                const uint32_t P = e->patterns[d.pattern].P;
                const uint64_t nwin = d.n_samples >= P ? d.n_samples - P + 1 : 0;
                uint64_t nseg = 1;
                const bool one_shot = d.chunk == 0 || d.chunk >= d.n_samples;
                if (one_shot && nwin > 2 * seg_min) {
                    const uint64_t share = (target_walkers * d.n_samples + total_samples - 1) / total_samples;
                    nseg = std::max<uint64_t>(1, std::min<uint64_t>(std::min<uint64_t>(share, nwin / seg_min), 60000));
                }
                const uint64_t unit = uint64_t(TMK_BLK) * 32;  // a walker reads block summaries 32 at a time
                uint64_t seg_len = nseg > 1 ? ((nwin + nseg - 1) / nseg + unit - 1) / unit * unit : std::max<uint64_t>(nwin, 1);
                nseg = nseg > 1 ? (nwin + seg_len - 1) / seg_len : 1;
                for (uint64_t k = 0; k < nseg; ++k) {
                    TmkSegItem it;
                    it.stream = s;
                    it.seg = uint32_t(k);
                    it.first_item = item_base[s];
                    it.n_seg = uint32_t(nseg);
                    it.q0 = k == 0 ? d.init_q : k * seg_len;
                    it.q1 = k + 1 < nseg ? (k + 1) * seg_len : ~0ull;
                    it.seg_len = seg_len;
                    items.push_back(it);
                }
            }
            item_base[n_streams] = uint32_t(items.size());
            seg_mode = items.size() > n_streams;
        }
    }
    const uint32_t n_items = seg_mode ? uint32_t(items.size()) : n_streams;
    TmkSegItem* d_items = nullptr;
    TmkSegState* d_segst = nullptr;
    uint32_t* d_first_valid = nullptr;
    if (seg_mode) {
        const size_t ib = sizeof(TmkSegItem) * n_items, sb = sizeof(TmkSegState) * n_items;
        CK(e->d_seg.reserve(ib + sb + sizeof(uint32_t) * n_items + 64));
        d_items = static_cast<TmkSegItem*>(e->d_seg.p);
        d_segst = reinterpret_cast<TmkSegState*>(static_cast<char*>(e->d_seg.p) + ib);
        d_first_valid = reinterpret_cast<uint32_t*>(static_cast<char*>(e->d_seg.p) + ib + sb);
        CK(cudaMemcpyAsync(d_items, items.data(), ib, cudaMemcpyHostToDevice, e->stream));
        CK(cudaMemsetAsync(d_segst, 0, sb, e->stream));
    }
    CK(e->d_cnt.reserve(sizeof(uint32_t) * n_items * (seg_mode ? 2 : 1)));
    CK(e->d_base.reserve(sizeof(uint64_t) * (n_items + 1)));
    CK(e->d_next.reserve(sizeof(unsigned int) * (2 * chunks.size() + 1)));
    // cluster pass (walk.cu): candidate entries + speculative verdicts, per chunk its own list region / counter
    // (short polls of carried-state links walk faster without it: the cluster pass and the walk of a poll are two dependent
    //  ~20 us single-warp frame decodes, one of which goes away — 256 links 0.131 -> 0.118 ms, one link 0.091 -> 0.075 ms)
    uint64_t stream_samples = 0;  // (not total_samples: the links pad their working buffer)
    for (uint32_t s = 0; s < n_streams; ++s) stream_samples += streams[s].n_samples;
    const bool short_polls = link_init != nullptr && stream_samples < 16384ull * n_streams;
    const bool use_fast = b.n_blocks > 0 && b.n_blocks < (1ull << 31) && !short_polls && !getenv("TMK_NO_CLUSTER");
    // (the walk of a chunk follows its cluster pass: running them side by side — the walk waiting for blocks the
    //  cluster pass had not reached — measured slower in round 1, 16.2 ms against 12.0 per step, and was removed)
    const bool serial_cluster = true;
    const uint64_t entry_cap = std::min<uint64_t>((1ull << 26) - 1, b.n_blocks / 2 + 4096);
    if (use_fast) {
        CK(e->d_cblk.reserve(sizeof(uint32_t) * b.n_blocks));
        CK(e->d_blk_ptr.reserve(sizeof(uint2) * b.n_blocks));  // (entry pointer, gain) per block with candidates
        CK(e->d_entries.reserve(sizeof(TmkCandEntry) * entry_cap));
        CK(e->d_pool.reserve(sizeof(unsigned int) * (2 + 2 * chunks.size())));
    }

    const TmkStreamDev* ds = static_cast<const TmkStreamDev*>(e->d_streams.p);
    const TmkPattern* dp = static_cast<const TmkPattern*>(e->d_patterns.p);
    float4* dbi = static_cast<float4*>(e->d_blk_info.p);
    uint32_t* dcb = static_cast<uint32_t*>(e->d_cand.p);
    TmkCounters* dc = static_cast<TmkCounters*>(e->d_counters.p);
    unsigned int* dnext = static_cast<unsigned int*>(e->d_next.p);
    unsigned int* dpool = use_fast ? static_cast<unsigned int*>(e->d_pool.p) : nullptr;

    if (reinterpret_cast<uintptr_t>(dx) % 16 != 0) return fail(e, TMK_E_INVALID, "sample buffer must be 16-byte aligned");
    CK(e->d_tile_dir.reserve(sizeof(uint32_t) * std::max<uint64_t>(1, b.n_tiles)));
    CK(e->d_tile_desc.reserve(tmk_scan_tile_desc_bytes(uint32_t(b.n_tiles)) + 64));
    uint32_t* dtd = static_cast<uint32_t*>(e->d_tile_dir.p);

    // record / payload scratch on the device: start from a guess, rerun the (cheap) walk if the
    // counts say it was too small and the caller's buffers are larger.
    const bool recs_dev = rec_cap && is_device_ptr(recs), pay_dev = payload_cap && is_device_ptr(payload);
    // (the walk's warps reserve record slots 32 at a time: the scratch holds kSlotSlack slots more than the records
    //  it is meant for, and the counter the walk returns counts reserved slots; the exact number is the sum of the
    //  per-stream counts)
    const uint64_t kSlotSlack = 32ull * 148 * 16 + 64;
    uint64_t dev_rec_cap = std::min<uint64_t>(rec_cap, std::max<uint64_t>(4096, total_samples / 192 + 2ull * n_streams)) + kSlotSlack;
    uint64_t dev_pay_cap = pay_dev ? payload_cap : std::min<uint64_t>(payload_cap, std::max<uint64_t>(65536, total_samples / 24));
    CK(e->d_recs_tmp.reserve(sizeof(tmk_frame_rec) * std::max<uint64_t>(1, dev_rec_cap)));
    CK(e->d_rec_ord.reserve(sizeof(uint4) * std::max<uint64_t>(1, dev_rec_cap)));
    // Small batches (the links' polls, single captures): counters, exit states, ordered records and payload are gathered in ONE
    // device block and reach pinned host memory with ONE copy before ONE synchronisation. The general path below needs the
    // counts on the host before it can size its copies: two synchronisations and five copies into pageable memory, each of
    // which the host waits for — ~100 us of a 260 us poll of 256 links.
    struct SmallHdr {
        TmkCounters hc;
        uint64_t n_exact;
        unsigned int next[4];
        unsigned int pool[4];
    };
    const uint64_t r_spec = dev_rec_cap - kSlotSlack;  // records the block has room for (= the caller's capacity, or the guess)
    const bool small = chunks.size() == 1 && !seg_mode && !recs_dev && !pay_dev && r_spec <= 65536 && dev_pay_cap <= (4u << 20) &&
                       n_streams <= 16384 && !getenv("TMK_NO_SMALL_PATH");
    // what the first copy brings over: the header, the exit states, and as many records / payload bytes as a poll usually has
    const uint64_t c_recs = std::min<uint64_t>(r_spec, 512 + n_streams / 2), c_pay = std::min<uint64_t>(dev_pay_cap, 32768 + 40ull * n_streams);
    const size_t sm_link = (sizeof(SmallHdr) + 15) & ~size_t(15);
    const size_t sm_recs = (sm_link + (link_final ? sizeof(TmkLinkState) * n_streams : 0) + 15) & ~size_t(15);
    // block: [header | exit states | first c_recs records | payload | the other records] — header .. first payload bytes is ONE copy
    const size_t sm_pay = (sm_recs + sizeof(tmk_frame_rec) * c_recs + 15) & ~size_t(15);
    const size_t sm_rest = (sm_pay + dev_pay_cap + 15) & ~size_t(15);
    const size_t sm_bytes = sm_rest + sizeof(tmk_frame_rec) * (r_spec - c_recs);
    char* sm = nullptr;
    if (small) {
        CK(e->d_small.reserve(sm_bytes));
        if (e->h_small_cap < sm_bytes) {
            if (e->h_small) cudaFreeHost(e->h_small);
            e->h_small = nullptr;
            e->h_small_cap = 0;
            CK(cudaHostAlloc(&e->h_small, sm_bytes + sm_bytes / 4, cudaHostAllocDefault));
            e->h_small_cap = sm_bytes + sm_bytes / 4;
        }
        sm = static_cast<char*>(e->d_small.p);
        SmallHdr* hdr = reinterpret_cast<SmallHdr*>(sm);
        dc = &hdr->hc;
        dnext = hdr->next;
        if (use_fast) dpool = hdr->pool;
        if (link_final) d_link = reinterpret_cast<TmkLinkState*>(sm + sm_link);
    } else if (!pay_dev) {
        CK(e->d_payload.reserve(std::max<uint64_t>(1, dev_pay_cap)));
    }
    uint8_t* dpay_cur = pay_dev ? payload : small ? reinterpret_cast<uint8_t*>(sm + sm_pay) : static_cast<uint8_t*>(e->d_payload.p);
    uint64_t spec_cap = std::min<uint64_t>(0x7fffffffull, 2 * dev_rec_cap + 1024);
    if (use_fast) CK(e->d_spec.reserve(sizeof(tmk_frame_rec) * spec_cap));
    auto fast_of = [&](size_t g, const Chunk& ch) {
        TmkFast f;
        std::memset(&f, 0, sizeof f);
        if (!use_fast) return f;
        unsigned int* pool = dpool;
        f.cblk_list = static_cast<uint32_t*>(e->d_cblk.p) + uint64_t(ch.t0) * TMK_TILE_BLKS;
        f.cblk_count = pool + 2 + g;
        f.item_next = pool + 2 + chunks.size() + g;
        f.blk_res = static_cast<uint2*>(e->d_blk_ptr.p);
        f.entries = static_cast<TmkCandEntry*>(e->d_entries.p);
        f.entry_cap = uint32_t(entry_cap);
        f.spec = static_cast<tmk_frame_rec*>(e->d_spec.p);
        f.spec_cap = uint32_t(spec_cap);
        f.pool = pool;
        return f;
    };

    if (trace) tr[1] = us_since(t_in);
    CK(cudaEventRecord(e->ev[1], e->stream));
#ifdef TMK_SCAN_ARM_A
    const bool fuse_dir = false;
#else
    const bool fuse_dir = chunks.size() == 1 && (b.kinds & 3u) != 0 && b.n_tiles > 0;  // the scan's descriptor kernel writes the directory too
#endif
    if (!fuse_dir) e->launches += tmk_launch_tile_dir(ds, n_streams, uint32_t(b.n_tiles), dtd, e->stream);
#ifdef TMK_SCAN_ARM_A
    CK(cudaMemsetAsync(dcb, 0, sizeof(uint32_t) * TMK_BLK_WORDS * b.n_blocks, e->stream));  // (the CUDA-core scan ORs its bits in)
#endif
    if (small) {
        CK(cudaMemsetAsync(sm, 0, sizeof(SmallHdr), e->stream));  // counters, work counters and list heads in one go
    } else {
        CK(cudaMemsetAsync(dc, 0, sizeof(TmkCounters), e->stream));
        CK(cudaMemsetAsync(dnext, 0, sizeof(unsigned int) * (2 * chunks.size() + 1), e->stream));
        if (use_fast) CK(cudaMemsetAsync(dpool, 0, sizeof(unsigned int) * (2 + 2 * chunks.size()), e->stream));
    }
    if (host_samples) {
        // the copy stream may only overwrite the staging buffer once earlier work on it is done
        CK(cudaEventRecord(e->cev[0], e->stream));
        CK(cudaStreamWaitEvent(e->copy_stream, e->cev[0], 0));
    }
    for (size_t g = 0; g < chunks.size(); ++g) {
        const Chunk& ch = chunks[g];
        if (host_samples) {
            if (pcm16)
                CK(cudaMemcpyAsync(const_cast<int16_t*>(dpcm) + ch.lo, pcm + ch.lo, (ch.hi - ch.lo) * sizeof(int16_t),
                                   cudaMemcpyHostToDevice, e->copy_stream));
            else
                CK(cudaMemcpyAsync(const_cast<float*>(dx) + ch.lo, samples + ch.lo, (ch.hi - ch.lo) * sizeof(float),
                                   cudaMemcpyHostToDevice, e->copy_stream));
            cudaEvent_t ce = e->cev[1 + (g & 1)];
            CK(cudaEventRecord(ce, e->copy_stream));
            CK(cudaStreamWaitEvent(e->stream, ce, 0));
        }
        if (pcm16) e->launches += tmk_launch_pcm16_to_f32(dpcm + ch.lo, const_cast<float*>(dx) + ch.lo, ch.hi - ch.lo, e->stream);
        const TmkFast fast = fast_of(g, ch);
        e->launches += tmk_launch_scan(dx, total_samples, ds, n_streams, dtd, ch.t0, ch.t1 - ch.t0, dp, b.kinds, dbi, dcb,
                                       const_cast<uint32_t*>(fast.cblk_list), fast.cblk_count, e->stream,
                                       static_cast<char*>(e->d_tile_desc.p) + tmk_scan_tile_desc_bytes(ch.t0) * (ch.t0 ? 1 : 0), nullptr, fuse_dir);
        if (chunks.size() == 1) CK(cudaEventRecord(e->ev[2], e->stream));
        if (trace) tr[2] = us_since(t_in);
        uint8_t* dpay = dpay_cur;
        // The walks / cluster passes of different chunks overlap each other and the later scans (high-priority
        // streams). Within a chunk the walk follows the cluster pass; they CAN run side by side (the cluster
        // pass publishes every block with a release store, a walk that reaches a block first waits for it,
        // bounded, or evaluates the block itself), but that measured slower (see serial_cluster above).
        // (one chunk — device-resident batches, the links' polls: everything in order on the engine's stream, no hand-overs
        //  between streams; a poll of 256 links is ~110 us of kernels and was ~170 us of API calls and stream hops)
        const bool one = chunks.size() == 1;
        cudaStream_t cs = one ? e->stream : e->walk_streams[(2 * g) % 4], ws = one ? e->stream : e->walk_streams[(2 * g + 1) % 4];
        if (!one) {
            cudaEvent_t se = e->cev[3 + (g & 1)];
            CK(cudaEventRecord(se, e->stream));
            CK(cudaStreamWaitEvent(cs, se, 0));
            CK(cudaStreamWaitEvent(ws, se, 0));
        }
        if (use_fast) {
            e->launches += tmk_launch_cluster(dx, ds, dtd, dp, dbi, dcb, fast, dpay, dev_pay_cap, dc, !serial_cluster, cs);
            if (one) {
                CK(cudaEventRecord(e->ev[5], cs));
                e->ev5_valid = true;
            }
            if (trace) tr[3] = us_since(t_in);
            if (serial_cluster && !one) {
                CK(cudaEventRecord(e->wev[0], cs));
                CK(cudaStreamWaitEvent(ws, e->wev[0], 0));
            }
        }
        if (seg_mode) {  // pass 1: every walker over its segment; pass 2: on to the meeting point with the next walkers
            for (uint32_t pass = 1; pass <= 2; ++pass)
                e->launches += tmk_launch_walk(dx, ds, item_base[ch.s0], item_base[ch.s1], dnext + (pass == 1 ? g : chunks.size() + 1 + g),
                                               fast, dp, dbi, dcb, static_cast<tmk_frame_rec*>(e->d_recs_tmp.p),
                                               static_cast<uint4*>(e->d_rec_ord.p), dev_rec_cap, dpay, dev_pay_cap,
                                               static_cast<uint32_t*>(e->d_cnt.p), dc, nullptr, ws, TmkSeg{d_items, d_segst, pass, 0});
        } else {
        e->launches += tmk_launch_walk(dx, ds, ch.s0, ch.s1, dnext + g, fast, dp, dbi, dcb,
                                       static_cast<tmk_frame_rec*>(e->d_recs_tmp.p), static_cast<uint4*>(e->d_rec_ord.p),
                                       dev_rec_cap, dpay, dev_pay_cap, static_cast<uint32_t*>(e->d_cnt.p), dc, d_link, ws);
        }
        CK(cudaGetLastError());
        if (trace) tr[4] = us_since(t_in);
    }
    if (chunks.size() != 1) CK(cudaEventRecord(e->ev[2], e->stream));  // end of the last scan
    if (chunks.size() != 1)
        for (int i = 0; i < 4; ++i) {
            CK(cudaEventRecord(e->wev[i], e->walk_streams[i]));
            CK(cudaStreamWaitEvent(e->stream, e->wev[i], 0));
        }
    TmkCounters hc;
    uint64_t n_exact = 0;  // records = sum of the per-stream counts
    bool small_tried = false;
    if (small) {
        CK(cudaEventRecord(e->ev[3], e->stream));
        SmallHdr* hdr = reinterpret_cast<SmallHdr*>(sm);
        e->launches += tmk_launch_order_small(static_cast<const uint32_t*>(e->d_cnt.p), static_cast<uint64_t*>(e->d_base.p), n_streams,
                                              static_cast<const tmk_frame_rec*>(e->d_recs_tmp.p), static_cast<const uint4*>(e->d_rec_ord.p), dc,
                                              dev_rec_cap, use_fast ? static_cast<const tmk_frame_rec*>(e->d_spec.p) : nullptr,
                                              reinterpret_cast<tmk_frame_rec*>(sm + sm_recs), c_recs, reinterpret_cast<tmk_frame_rec*>(sm + sm_rest),
                                              r_spec, &hdr->n_exact, e->stream);
        CK(cudaGetLastError());
        CK(cudaEventRecord(e->ev[4], e->stream));
        char* hs = static_cast<char*>(e->h_small);
        CK(cudaMemcpyAsync(hs, sm, sm_pay + c_pay, cudaMemcpyDeviceToHost, e->stream));
        const double t_enq = trace ? us_since(t_in) : 0.0;
        CK(cudaStreamSynchronize(e->stream));  // (polling cudaStreamQuery instead measured the same: the runtime spins already)
        if (trace)
            fprintf(stderr, "tmk small batch (us from entry): prepared %.1f, reserved %.1f, scan %.1f, cluster %.1f, walk %.1f, all enqueued %.1f, synchronised %.1f\n",
                    tr[0], tr[1], tr[2], tr[3], tr[4], t_enq, us_since(t_in));
        const SmallHdr* hh = reinterpret_cast<const SmallHdr*>(hs);
        hc = hh->hc;
        n_exact = hh->n_exact;
        small_tried = true;
        if (hc.n_recs <= dev_rec_cap && hc.n_payload <= dev_pay_cap && n_exact <= r_spec) {  // (r_spec <= rec_cap, dev_pay_cap <= payload_cap)
            if (n_exact > c_recs || hc.n_payload > c_pay) {  // a busy poll: the rest, now that the counts are known
                if (n_exact > c_recs)
                    CK(cudaMemcpyAsync(hs + sm_rest, sm + sm_rest, sizeof(tmk_frame_rec) * (n_exact - c_recs), cudaMemcpyDeviceToHost, e->stream));
                if (hc.n_payload > c_pay)
                    CK(cudaMemcpyAsync(hs + sm_pay + c_pay, sm + sm_pay + c_pay, hc.n_payload - c_pay, cudaMemcpyDeviceToHost, e->stream));
                CK(cudaStreamSynchronize(e->stream));
            }
            if (n_exact) std::memcpy(recs, hs + sm_recs, sizeof(tmk_frame_rec) * std::min<uint64_t>(n_exact, c_recs));
            if (n_exact > c_recs) std::memcpy(recs + c_recs, hs + sm_rest, sizeof(tmk_frame_rec) * (n_exact - c_recs));
            if (hc.n_payload) std::memcpy(payload, hs + sm_pay, hc.n_payload);
            if (link_final) std::memcpy(link_final, hs + sm_link, sizeof(TmkLinkState) * n_streams);
            *n_recs = n_exact;
            *payload_len = hc.n_payload;
            for (int i = 0; i < TMK_STAT_COUNT; ++i) e->stats[i] += hc.stats[i];
            for (int i = 0; i < 8; ++i) e->dbg[i] = hc.dbg[i];
            e->stage_pending = true;  // (stage times: on demand, tmk_flush_stage_ms)
            return TMK_OK;
        }
        // the guesses were too small: the general path takes over with the counts it now has
    }
    if (small_tried) {
    } else if (seg_mode) {
        // which records of which walker count (meeting points), their verdict tallies, then the counts per walker
        uint32_t* d_nvalid = static_cast<uint32_t*>(e->d_cnt.p) + n_items;
        e->launches += tmk_launch_seg_resolve(d_items, d_segst, n_items, d_first_valid, d_nvalid,
                                              static_cast<const tmk_frame_rec*>(e->d_recs_tmp.p), static_cast<const uint4*>(e->d_rec_ord.p),
                                              dev_rec_cap, dev_rec_cap, use_fast ? static_cast<const tmk_frame_rec*>(e->d_spec.p) : nullptr,
                                              dc, e->stream);
        e->launches += tmk_launch_excl_scan(d_nvalid, static_cast<uint64_t*>(e->d_base.p), n_items, e->stream);
        CK(cudaMemcpyAsync(&n_exact, static_cast<uint64_t*>(e->d_base.p) + n_items, sizeof n_exact, cudaMemcpyDeviceToHost, e->stream));
    } else {
    e->launches += tmk_launch_excl_scan(static_cast<const uint32_t*>(e->d_cnt.p), static_cast<uint64_t*>(e->d_base.p), n_streams, e->stream);
    CK(cudaMemcpyAsync(&n_exact, static_cast<uint64_t*>(e->d_base.p) + n_streams, sizeof n_exact, cudaMemcpyDeviceToHost, e->stream));
    }
    if (!small_tried) {
        CK(cudaMemcpyAsync(&hc, dc, sizeof hc, cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
    }
    for (int attempt = 0; attempt < 2; ++attempt) {
        // (the counters of the first pass include what is dropped later: payload bytes of speculative decodes that no
        //  walk linked, records and bytes of segment walkers before their meeting point — before a caller's exactly
        //  sized buffer is called too small, the plain walk counts again)
        const bool dead = attempt == 0 && (seg_mode || use_fast);
        const bool rec_short = hc.n_recs > dev_rec_cap && (dev_rec_cap < rec_cap + kSlotSlack || (dead && seg_mode));
        const bool pay_short = hc.n_payload > dev_pay_cap && (dev_pay_cap < payload_cap || dead);
        if (!rec_short && !pay_short) break;
        // the guess was too small: grow the scratch and walk every stream again (everything is resident now)
        seg_mode = false;  // (one walker per capture in this rare second pass)
        if (rec_short) dev_rec_cap = std::min<uint64_t>(rec_cap, hc.n_recs) + kSlotSlack;
        if (pay_short) dev_pay_cap = std::min<uint64_t>(payload_cap, hc.n_payload);
        CK(e->d_recs_tmp.reserve(sizeof(tmk_frame_rec) * std::max<uint64_t>(1, dev_rec_cap)));
        CK(e->d_rec_ord.reserve(sizeof(uint4) * std::max<uint64_t>(1, dev_rec_cap)));
        if (!pay_dev) {
            CK(e->d_payload.reserve(std::max<uint64_t>(1, dev_pay_cap)));
            dpay_cur = static_cast<uint8_t*>(e->d_payload.p);
        }
        uint8_t* dpay = dpay_cur;
        CK(cudaMemsetAsync(dc, 0, sizeof(TmkCounters), e->stream));
        CK(cudaMemsetAsync(dnext + chunks.size(), 0, sizeof(unsigned int), e->stream));
        // the speculative verdicts of the cluster pass point into the payload scratch that was just re-sized:
        // this (rare) second pass lets the walk evaluate everything itself
        TmkFast nofast;
        std::memset(&nofast, 0, sizeof nofast);
        e->launches += tmk_launch_walk(dx, ds, 0, n_streams, dnext + chunks.size(), nofast, dp, dbi, dcb,
                                       static_cast<tmk_frame_rec*>(e->d_recs_tmp.p), static_cast<uint4*>(e->d_rec_ord.p),
                                       dev_rec_cap, dpay, dev_pay_cap, static_cast<uint32_t*>(e->d_cnt.p), dc, d_link, e->stream);
        CK(cudaGetLastError());
        e->launches += tmk_launch_excl_scan(static_cast<const uint32_t*>(e->d_cnt.p), static_cast<uint64_t*>(e->d_base.p), n_streams, e->stream);
        CK(cudaMemcpyAsync(&n_exact, static_cast<uint64_t*>(e->d_base.p) + n_streams, sizeof n_exact, cudaMemcpyDeviceToHost, e->stream));
        CK(cudaMemcpyAsync(&hc, dc, sizeof hc, cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
    }
    if (link_final)
        CK(cudaMemcpyAsync(link_final, d_link, sizeof(TmkLinkState) * n_streams, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaEventRecord(e->ev[3], e->stream));
    *n_recs = n_exact;
    *payload_len = hc.n_payload;
    for (int i = 0; i < TMK_STAT_COUNT; ++i) e->stats[i] += hc.stats[i];
    for (int i = 0; i < 8; ++i) e->dbg[i] = hc.dbg[i];
    if (chunks.size() == 1) {  // [4]: device time of the cluster pass in microseconds (part of stage [1])
        float cms = 0.f;
        if (e->ev5_valid && cudaEventElapsedTime(&cms, e->ev[2], e->ev[5]) == cudaSuccess) e->dbg[4] = uint64_t(cms * 1000.f);
        cudaGetLastError();  // (an event that was never recorded must not leave an error behind)
    }
    const bool overflow = n_exact > rec_cap || hc.n_recs > dev_rec_cap || hc.n_payload > payload_cap;
    if (!overflow && n_exact) {
        tmk_frame_rec* dout = recs;
        if (!recs_dev) {
            CK(e->d_recs.reserve(sizeof(tmk_frame_rec) * n_exact));
            dout = static_cast<tmk_frame_rec*>(e->d_recs.p);
        }
        if (seg_mode)
            e->launches += tmk_launch_seg_order(static_cast<const tmk_frame_rec*>(e->d_recs_tmp.p),
                                                static_cast<const uint4*>(e->d_rec_ord.p), hc.n_recs,
                                                static_cast<const uint64_t*>(e->d_base.p),
                                                use_fast ? static_cast<const tmk_frame_rec*>(e->d_spec.p) : nullptr, d_items, d_first_valid,
                                                dout, e->stream);
        else
        e->launches += tmk_launch_order(static_cast<const tmk_frame_rec*>(e->d_recs_tmp.p),
                                        static_cast<const uint4*>(e->d_rec_ord.p), hc.n_recs,
                                        static_cast<const uint32_t*>(e->d_cnt.p), static_cast<uint64_t*>(e->d_base.p),
                                        n_streams, use_fast ? static_cast<const tmk_frame_rec*>(e->d_spec.p) : nullptr, dout, e->stream);
        CK(cudaGetLastError());
        CK(cudaEventRecord(e->ev[4], e->stream));
        if (!recs_dev)
            CK(cudaMemcpyAsync(recs, dout, sizeof(tmk_frame_rec) * n_exact, cudaMemcpyDeviceToHost, e->stream));
        if (!pay_dev && hc.n_payload)
            CK(cudaMemcpyAsync(payload, dpay_cur, hc.n_payload, cudaMemcpyDeviceToHost, e->stream));
    } else {
        CK(cudaEventRecord(e->ev[4], e->stream));
    }
    CK(cudaStreamSynchronize(e->stream));
    cudaEventElapsedTime(&e->stage_ms[3], e->ev[0], e->ev[1]);
    cudaEventElapsedTime(&e->stage_ms[0], e->ev[1], e->ev[2]);
    cudaEventElapsedTime(&e->stage_ms[1], e->ev[2], e->ev[3]);
    cudaEventElapsedTime(&e->stage_ms[2], e->ev[3], e->ev[4]);
    if (overflow) return fail(e, TMK_E_CAPACITY, "tmk_decode_batch: record or payload buffer too small");
    return TMK_OK;
}

int tmk_decode_batch(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                     uint32_t n_streams, tmk_frame_rec* recs, uint64_t rec_cap, uint64_t* n_recs, uint8_t* payload,
                     uint64_t payload_cap, uint64_t* payload_len) {
    return decode_batch_impl(e, samples, false, total_samples, streams, n_streams, recs, rec_cap, n_recs, payload, payload_cap,
                             payload_len);
}

int tmk_decode_batch_pcm16(tmk_engine* e, const int16_t* pcm, uint64_t total_samples, const tmk_stream* streams,
                           uint32_t n_streams, tmk_frame_rec* recs, uint64_t rec_cap, uint64_t* n_recs, uint8_t* payload,
                           uint64_t payload_cap, uint64_t* payload_len) {
    return decode_batch_impl(e, pcm, true, total_samples, streams, n_streams, recs, rec_cap, n_recs, payload, payload_cap,
                             payload_len);
}

int tmk_pcm16_encode(tmk_engine* e, const float* samples, uint64_t n, int16_t* pcm_out) {
    if (!e || ((!samples || !pcm_out) && n)) return fail(e, TMK_E_INVALID, "tmk_pcm16_encode: null argument");
    if (n == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    const void* d_in = nullptr;
    int rc = stage_in(e, e->d_samples, samples, n * sizeof(float), &d_in);
    if (rc != TMK_OK) return rc;
    int16_t* d_out = pcm_out;
    const bool dev_out = is_device_ptr(pcm_out);
    if (!dev_out) {
        CK(e->d_pcm.reserve(n * sizeof(int16_t)));
        d_out = static_cast<int16_t*>(e->d_pcm.p);
    }
    if (reinterpret_cast<uintptr_t>(d_in) % 16 != 0 || reinterpret_cast<uintptr_t>(d_out) % 8 != 0)
        return fail(e, TMK_E_INVALID, "tmk_pcm16_encode: device pointers must be aligned (samples 16 bytes, pcm 8 bytes)");
    e->launches += tmk_launch_pcm16_encode(static_cast<const float*>(d_in), d_out, n, e->stream);
    CK(cudaGetLastError());
    if (!dev_out) CK(cudaMemcpyAsync(pcm_out, d_out, n * sizeof(int16_t), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_pcm16_to_f32(tmk_engine* e, const int16_t* pcm, uint64_t n, float* samples_out) {
    if (!e || ((!pcm || !samples_out) && n)) return fail(e, TMK_E_INVALID, "tmk_pcm16_to_f32: null argument");
    if (n == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    const void* d_in = nullptr;
    int rc = stage_in(e, e->d_pcm, pcm, n * sizeof(int16_t), &d_in);
    if (rc != TMK_OK) return rc;
    float* d_out = samples_out;
    const bool dev_out = is_device_ptr(samples_out);
    if (!dev_out) {
        CK(e->d_samples.reserve(n * sizeof(float)));
        d_out = static_cast<float*>(e->d_samples.p);
    }
    if (reinterpret_cast<uintptr_t>(d_in) % 8 != 0 || reinterpret_cast<uintptr_t>(d_out) % 16 != 0)
        return fail(e, TMK_E_INVALID, "tmk_pcm16_to_f32: device pointers must be aligned (pcm 8 bytes, samples 16 bytes)");
    e->launches += tmk_launch_pcm16_to_f32(static_cast<const int16_t*>(d_in), d_out, n, e->stream);
    CK(cudaGetLastError());
    if (!dev_out) CK(cudaMemcpyAsync(samples_out, d_out, n * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_correlate(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                  uint32_t n_streams, float* corr_out) {
    if (!e || !samples || !streams || !corr_out) return fail(e, TMK_E_INVALID, "tmk_correlate: null argument");
    CK(cudaSetDevice(e->device));
    Batch b;
    int rc = prepare_batch(e, total_samples, streams, n_streams, &b);
    if (rc != TMK_OK) return rc;
    CK(e->d_streams.reserve(sizeof(TmkStreamDev) * n_streams));
    CK(cudaMemcpyAsync(e->d_streams.p, b.sd.data(), sizeof(TmkStreamDev) * n_streams, cudaMemcpyHostToDevice, e->stream));
    const void* d_samples = nullptr;
    rc = stage_in(e, e->d_samples, samples, total_samples * sizeof(float), &d_samples);
    if (rc != TMK_OK) return rc;
    float* d_out = corr_out;
    const bool dev_out = is_device_ptr(corr_out);
    if (!dev_out) {
        CK(e->d_misc2.reserve(total_samples * sizeof(float)));
        d_out = static_cast<float*>(e->d_misc2.p);
        CK(cudaMemsetAsync(d_out, 0, total_samples * sizeof(float), e->stream));
    }
    e->launches += tmk_launch_correlate(static_cast<const float*>(d_samples), static_cast<const TmkStreamDev*>(e->d_streams.p),
                                        n_streams, static_cast<const TmkPattern*>(e->d_patterns.p), d_out, e->stream);
    CK(cudaGetLastError());
    if (!dev_out) CK(cudaMemcpyAsync(corr_out, d_out, total_samples * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_scan_probe(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                   uint32_t n_streams, float* blk_info, uint32_t* cand_bits, uint64_t block_cap, uint64_t* n_blocks,
                   uint64_t* blk_base) {
#ifndef TMK_SCAN_ARM_A
    static_assert(TMK_SCAN_BLOCK == TMK_BLK, "header and kernels disagree on the scan block size");
#endif
    if (!e || !samples || !streams || !blk_info || !cand_bits || !n_blocks || !blk_base)
        return fail(e, TMK_E_INVALID, "tmk_scan_probe: null argument");
    CK(cudaSetDevice(e->device));
    Batch b;
    int rc = prepare_batch(e, total_samples, streams, n_streams, &b);
    if (rc != TMK_OK) return rc;
    *n_blocks = b.n_blocks;
    for (uint32_t s = 0; s < n_streams; ++s) blk_base[s] = b.sd[s].blk_base;
    if (b.n_blocks > block_cap) return fail(e, TMK_E_CAPACITY, "tmk_scan_probe: block buffers too small");
    if (b.n_blocks == 0) return TMK_OK;
    CK(e->d_streams.reserve(sizeof(TmkStreamDev) * n_streams));
    CK(cudaMemcpyAsync(e->d_streams.p, b.sd.data(), sizeof(TmkStreamDev) * n_streams, cudaMemcpyHostToDevice, e->stream));
    const void* d_samples = nullptr;
    rc = stage_in(e, e->d_samples, samples, total_samples * sizeof(float), &d_samples);
    if (rc != TMK_OK) return rc;
    CK(e->d_blk_info.reserve(sizeof(float4) * b.n_blocks));
    CK(e->d_cand.reserve(sizeof(uint32_t) * TMK_BLK_WORDS * b.n_blocks));
    CK(e->d_tile_dir.reserve(sizeof(uint32_t) * b.n_tiles));
    CK(e->d_tile_desc.reserve(tmk_scan_tile_desc_bytes(uint32_t(b.n_tiles)) + 64));
    const TmkStreamDev* ds = static_cast<const TmkStreamDev*>(e->d_streams.p);
    uint32_t* dtd = static_cast<uint32_t*>(e->d_tile_dir.p);
    CK(cudaMemsetAsync(e->d_blk_info.p, 0, sizeof(float4) * b.n_blocks, e->stream));
    CK(cudaMemsetAsync(e->d_cand.p, 0, sizeof(uint32_t) * TMK_BLK_WORDS * b.n_blocks, e->stream));
    e->launches += tmk_launch_tile_dir(ds, n_streams, uint32_t(b.n_tiles), dtd, e->stream);
    e->launches += tmk_launch_scan(static_cast<const float*>(d_samples), total_samples, ds, n_streams, dtd, 0,
                                   uint32_t(b.n_tiles), static_cast<const TmkPattern*>(e->d_patterns.p), b.kinds,
                                   static_cast<float4*>(e->d_blk_info.p), static_cast<uint32_t*>(e->d_cand.p), nullptr, nullptr,
                                   e->stream, e->d_tile_desc.p, e->scan_dbg_dot);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(blk_info, e->d_blk_info.p, sizeof(float4) * b.n_blocks, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(cand_bits, e->d_cand.p, sizeof(uint32_t) * TMK_BLK_WORDS * b.n_blocks, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

uint32_t tmk_scan_block(void) { return TMK_BLK; }

// Diagnostic: the correlation numerators D(w) = sum_t x[w + t] p[t] the scan filter computed (TF32 tensor-core
// products in the default build), dot_out[stream offset + w]; windows the filter does not cover stay 0.
int tmk_scan_dot_probe(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                       uint32_t n_streams, float* dot_out) {
    if (!e || !samples || !streams || !dot_out) return fail(e, TMK_E_INVALID, "tmk_scan_dot_probe: null argument");
    if (is_device_ptr(dot_out)) return fail(e, TMK_E_INVALID, "tmk_scan_dot_probe: host output only");
    CK(cudaSetDevice(e->device));
    Batch b;
    int rc = prepare_batch(e, total_samples, streams, n_streams, &b);
    if (rc != TMK_OK) return rc;
    std::vector<float> info(4 * std::max<uint64_t>(1, b.n_blocks));
    std::vector<uint32_t> bits(TMK_BLK_WORDS * std::max<uint64_t>(1, b.n_blocks));
    std::vector<uint64_t> base(n_streams);
    uint64_t nb = 0;
    CK(e->d_misc2.reserve(total_samples * sizeof(float)));
    CK(cudaMemsetAsync(e->d_misc2.p, 0, total_samples * sizeof(float), e->stream));
    e->scan_dbg_dot = static_cast<float*>(e->d_misc2.p);
    rc = tmk_scan_probe(e, samples, total_samples, streams, n_streams, info.data(), bits.data(), b.n_blocks, &nb, base.data());
    e->scan_dbg_dot = nullptr;
    if (rc != TMK_OK) return rc;
    CK(cudaMemcpyAsync(dot_out, e->d_misc2.p, total_samples * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_replay_probe(tmk_engine* e, const float* samples, uint64_t n_samples, const tmk_phy_config* cfg, uint64_t seed,
                     const uint64_t* windows, uint32_t n_windows, float* energy, uint64_t counts[3]) {
    if (!e || !samples || !cfg || (!windows && n_windows) || (!energy && n_windows))
        return fail(e, TMK_E_INVALID, "tmk_replay_probe: null argument");
    CK(cudaSetDevice(e->device));
    tmk_stream st;
    st.offset = 0;
    st.n_samples = n_samples;
    st.cfg = *cfg;
    st.cfg.chunk = 0;
    Batch b;
    int rc = prepare_batch(e, n_samples, &st, 1, &b);
    if (rc != TMK_OK) return rc;
    const uint64_t P = e->patterns[b.sd[0].pattern].P;
    uint64_t prev = seed;
    for (uint32_t i = 0; i < n_windows; ++i) {
        if (windows[i] < prev || windows[i] + P > n_samples)
            return fail(e, TMK_E_INVALID, "tmk_replay_probe: windows must ascend from the seed and lie inside the capture");
        prev = windows[i];
    }
    if (n_windows == 0) return TMK_OK;
    CK(e->d_streams.reserve(sizeof(TmkStreamDev)));
    CK(cudaMemcpyAsync(e->d_streams.p, b.sd.data(), sizeof(TmkStreamDev), cudaMemcpyHostToDevice, e->stream));
    const void* d_samples = nullptr;
    rc = stage_in(e, e->d_samples, samples, n_samples * sizeof(float), &d_samples);
    if (rc != TMK_OK) return rc;
    const size_t wbytes = sizeof(uint64_t) * n_windows, obytes = sizeof(float) * n_windows;
    CK(e->d_misc.reserve(wbytes + obytes + 64));
    char* dm = static_cast<char*>(e->d_misc.p);
    CK(cudaMemcpyAsync(dm, windows, wbytes, cudaMemcpyHostToDevice, e->stream));
    CK(e->d_misc2.reserve(64));
    e->launches += tmk_launch_replay_probe(static_cast<const float*>(d_samples), static_cast<const TmkStreamDev*>(e->d_streams.p),
                                           static_cast<const TmkPattern*>(e->d_patterns.p), seed,
                                           reinterpret_cast<const uint64_t*>(dm), n_windows,
                                           reinterpret_cast<float*>(dm + wbytes), static_cast<unsigned long long*>(e->d_misc2.p),
                                           e->stream);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(energy, dm + wbytes, obytes, cudaMemcpyDeviceToHost, e->stream));
    unsigned long long hc[3] = {0, 0, 0};
    CK(cudaMemcpyAsync(hc, e->d_misc2.p, sizeof hc, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    if (counts)
        for (int i = 0; i < 3; ++i) counts[i] = hc[i];
    return TMK_OK;
}

int tmk_channel_busy(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                     uint32_t n_streams, int8_t* busy) {
    if (!e || (!samples && total_samples) || !streams || !busy) return fail(e, TMK_E_INVALID, "tmk_channel_busy: null argument");
    CK(cudaSetDevice(e->device));
    std::vector<TmkStreamDev> sd(n_streams);
    for (uint32_t s = 0; s < n_streams; ++s) {
        if (streams[s].offset > total_samples || streams[s].n_samples > total_samples - streams[s].offset)
            return fail(e, TMK_E_INVALID, "tmk_channel_busy: stream exceeds the sample buffer");
        std::memset(&sd[s], 0, sizeof sd[s]);
        sd[s].offset = streams[s].offset;
        sd[s].n_samples = streams[s].n_samples;
    }
    CK(e->d_streams.reserve(sizeof(TmkStreamDev) * std::max<uint32_t>(1, n_streams)));
    CK(cudaMemcpyAsync(e->d_streams.p, sd.data(), sizeof(TmkStreamDev) * n_streams, cudaMemcpyHostToDevice, e->stream));
    const void* d_samples = nullptr;
    int rc = stage_in(e, e->d_samples, samples, total_samples * sizeof(float), &d_samples);
    if (rc != TMK_OK) return rc;
    int8_t* d_out = busy;
    const bool dev_out = is_device_ptr(busy);
    if (!dev_out) {
        CK(e->d_misc2.reserve(std::max<uint32_t>(1, n_streams)));
        d_out = static_cast<int8_t*>(e->d_misc2.p);
    }
    e->launches += tmk_launch_busy(static_cast<const float*>(d_samples), static_cast<const TmkStreamDev*>(e->d_streams.p),
                                   n_streams, d_out, e->stream);
    CK(cudaGetLastError());
    if (!dev_out && n_streams) CK(cudaMemcpyAsync(busy, d_out, n_streams, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_stats(tmk_engine* e, uint64_t counters[TMK_STAT_COUNT]) {
    if (!e || !counters) return TMK_E_INVALID;
    for (int i = 0; i < TMK_STAT_COUNT; ++i) counters[i] = e->stats[i];
    return TMK_OK;
}

int tmk_debug_counters(tmk_engine* e, uint64_t out[8]) {
    if (!e || !out) return TMK_E_INVALID;
    cudaSetDevice(e->device);
    tmk_flush_stage_ms(e);
    for (int i = 0; i < 8; ++i) out[i] = e->dbg[i];
    return TMK_OK;
}

int tmk_stats_reset(tmk_engine* e) {
    if (!e) return TMK_E_INVALID;
    for (int i = 0; i < TMK_STAT_COUNT; ++i) e->stats[i] = 0;
    return TMK_OK;
}

int tmk_device_alloc(tmk_engine* e, uint64_t bytes, void** out) {
    if (!e || !out) return TMK_E_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaMalloc(out, bytes ? bytes : 1));
    return TMK_OK;
}

int tmk_device_free(tmk_engine* e, void* p) {
    if (!e) return TMK_E_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaFree(p));
    return TMK_OK;
}

int tmk_host_alloc_pinned(tmk_engine* e, uint64_t bytes, void** out) {
    if (!e || !out) return TMK_E_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaMallocHost(out, bytes ? bytes : 1));
    return TMK_OK;
}

int tmk_host_free_pinned(tmk_engine* e, void* p) {
    if (!e) return TMK_E_INVALID;
    CK(cudaFreeHost(p));
    return TMK_OK;
}

int tmk_memcpy_h2d(tmk_engine* e, void* dst, const void* src, uint64_t bytes) {
    if (!e) return TMK_E_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_memcpy_d2h(tmk_engine* e, void* dst, const void* src, uint64_t bytes) {
    if (!e) return TMK_E_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_synchronize(tmk_engine* e) {
    if (!e) return TMK_E_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

void* tmk_engine_stream(tmk_engine* e) { return e ? static_cast<void*>(e->stream) : nullptr; }

int tmk_last_kernel_ms(tmk_engine* e, float* ms) {
    if (!e || !ms) return TMK_E_INVALID;
    cudaSetDevice(e->device);
    tmk_flush_stage_ms(e);
    *ms = e->stage_ms[0] + e->stage_ms[1] + e->stage_ms[2];
    return TMK_OK;
}

int tmk_last_stage_ms(tmk_engine* e, float ms[4]) {
    if (!e || !ms) return TMK_E_INVALID;
    cudaSetDevice(e->device);
    tmk_flush_stage_ms(e);
    for (int i = 0; i < 4; ++i) ms[i] = e->stage_ms[i];
    return TMK_OK;
}

int tmk_timer_start(tmk_engine* e) {
    if (!e) return TMK_E_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->stream));
    CK(cudaEventRecord(e->tev[0], e->stream));
    return TMK_OK;
}

int tmk_timer_stop(tmk_engine* e, float* ms) {
    if (!e || !ms) return TMK_E_INVALID;
    CK(cudaEventRecord(e->tev[1], e->stream));
    CK(cudaEventSynchronize(e->tev[1]));
    CK(cudaEventElapsedTime(ms, e->tev[0], e->tev[1]));
    return TMK_OK;
}

int tmk_device_pci_bus_id(tmk_engine* e, char* out, int cap) {
    if (!e || !out || cap < 16) return TMK_E_INVALID;
    CK(cudaDeviceGetPCIBusId(out, cap, e->device));
    for (char* c = out; *c; ++c)
        if (*c >= 'A' && *c <= 'Z') *c = char(*c - 'A' + 'a');
    return TMK_OK;
}

uint64_t tmk_launch_count(const tmk_engine* e) { return e ? e->launches : 0; }

int tmk_synth_batch_ex(tmk_engine* e, const tmk_phy_config* cfg, uint64_t seed, uint32_t n_streams, uint64_t n_per_stream,
                       const tmk_synth_params* prm, float* samples, tmk_frame_rec* truth, uint64_t truth_cap,
                       uint64_t* n_truth) {
    if (!e || !cfg || !prm || !samples || !n_truth || (!truth && truth_cap))
        return fail(e, TMK_E_INVALID, "tmk_synth_batch: null argument");
    if (!valid_cfg(*cfg) || n_per_stream % 4 != 0 || n_per_stream >= (1ull << 32))
        return fail(e, TMK_E_INVALID, "tmk_synth_batch: invalid config or n_per_stream (multiple of 4, < 2^32)");
    TmkSynthImp imp;
    imp.multipath = prm->multipath ? 1u : 0u;
    imp.drift_ppm = prm->drift_ppm;
    imp.gap_min = prm->gap_max > prm->gap_min ? prm->gap_min : 48u;
    imp.gap_span = prm->gap_max > prm->gap_min ? prm->gap_max - prm->gap_min : 4800u - 48u;
    imp.amp_min = prm->amp_min;
    if (!(std::fabs(imp.drift_ppm) <= 10000.0f) || imp.amp_min < 0.0f || imp.amp_min > 1.0f || (imp.multipath && imp.gap_min < 48u))
        return fail(e, TMK_E_INVALID, "tmk_synth_batch_ex: drift beyond +-10000 ppm, amp_min outside [0, 1] or gaps below 48 with multipath");
    const float snr_db = prm->snr_db, amplitude = prm->amplitude;
    CK(cudaSetDevice(e->device));
    uint32_t pi = 0;
    int rc = pattern_index(e, *cfg, &pi);
    if (rc != TMK_OK) return rc;
    const TmkPattern* dp = static_cast<const TmkPattern*>(e->d_patterns.p) + pi;
    const uint64_t total = uint64_t(n_streams) * n_per_stream;
    float* d_x = samples;
    const bool dev_out = is_device_ptr(samples);
    if (!dev_out) {
        CK(e->d_samples.reserve(total * sizeof(float)));
        d_x = static_cast<float*>(e->d_samples.p);
    }
    const float sigma = snr_db >= 100.0f ? 0.0f : amplitude * std::pow(10.0f, -snr_db / 20.0f);
    e->launches += tmk_launch_synth_noise(seed, total, sigma, n_per_stream, imp, d_x, e->stream);
    CK(e->d_cnt.reserve(sizeof(uint32_t) * std::max<uint32_t>(1, n_streams)));
    CK(e->d_base.reserve(sizeof(uint64_t) * (n_streams + 1)));
    uint32_t* d_cnt = static_cast<uint32_t*>(e->d_cnt.p);
    uint64_t* d_base = static_cast<uint64_t*>(e->d_base.p);
    e->launches += tmk_launch_synth_plan(dp, cfg->local_addr, seed, n_streams, n_per_stream, imp, d_cnt, nullptr, nullptr, 0, e->stream);
    e->launches += tmk_launch_excl_scan(d_cnt, d_base, n_streams, e->stream);
    uint64_t n_frames = 0;
    CK(cudaMemcpyAsync(&n_frames, d_base + n_streams, sizeof n_frames, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    *n_truth = n_frames;
    CK(e->d_recs_tmp.reserve(sizeof(tmk_frame_rec) * std::max<uint64_t>(1, n_frames)));
    tmk_frame_rec* d_fr = static_cast<tmk_frame_rec*>(e->d_recs_tmp.p);
    e->launches += tmk_launch_synth_plan(dp, cfg->local_addr, seed, n_streams, n_per_stream, imp, d_cnt, d_base, d_fr, n_frames, e->stream);
    e->launches += tmk_launch_synth_frames(dp, seed, n_per_stream, amplitude, imp, d_fr, n_frames, d_x, e->stream);
    CK(cudaGetLastError());
    if (!dev_out) CK(cudaMemcpyAsync(samples, d_x, total * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    if (truth && n_frames) {
        const uint64_t n = std::min<uint64_t>(n_frames, truth_cap);
        CK(cudaMemcpyAsync(truth, d_fr, sizeof(tmk_frame_rec) * n,
                           is_device_ptr(truth) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, e->stream));
    }
    CK(cudaStreamSynchronize(e->stream));
    if (truth && n_frames > truth_cap) return fail(e, TMK_E_CAPACITY, "tmk_synth_batch: truth buffer too small");
    return TMK_OK;
}

int tmk_synth_batch(tmk_engine* e, const tmk_phy_config* cfg, uint64_t seed, uint32_t n_streams, uint64_t n_per_stream,
                    float snr_db, float amplitude, float* samples, tmk_frame_rec* truth, uint64_t truth_cap,
                    uint64_t* n_truth) {
    tmk_synth_params prm;
    std::memset(&prm, 0, sizeof prm);
    prm.snr_db = snr_db;
    prm.amplitude = amplitude;
    return tmk_synth_batch_ex(e, cfg, seed, n_streams, n_per_stream, &prm, samples, truth, truth_cap, n_truth);
}

}  // extern "C"
