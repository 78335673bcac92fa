// links.cu — carried-state streaming decode: n PhyDecoder instances (one per acoustic link) fed with
// process_samples calls, all links of a call decoded in ONE batch.
//
// The reference's PhyDecoder (src/phy/decoder.rs:16-33) carries between process_samples calls:
//   sample_buffer (drained to P-1 samples before buffer_offset, :90-109), buffer_offset, and the state
//   Searching | Decoding(frame_start).
// One call appends the new samples and runs search / decode_frame over the whole buffer until one of
// them needs more data (:78-88). That is exactly a one-shot decode of [retained samples | new samples]
// entered at buffer_offset in the carried state, so a call here is:
//   1. assemble_kernel: working buffer of every link = its retained tail ++ its new samples (HBM to HBM);
//   2. tmk_decode_batch over the working buffers (scan, cluster pass, walk) with the entry state of every
//      link (TmkStreamDev::init_q / init_fs) and the exit state written by the walk (TmkLinkState);
//   3. the drain rule (:90-109) applied per link: the retained tail STAYS where it is in the working buffer (two working
//      buffers take turns), the next call assembles from there — no second copy, no second synchronisation.
// Records carry positions counted from the first sample ever fed to the link; a frame still being
// gathered (TRUNCATED in the offline engine) is state, not a record.
#include <algorithm>
#include <cstring>
#include <vector>

#include "engine.hpp"

struct tmk_links {
    tmk_engine* e = nullptr;
    uint32_t n = 0;
    std::vector<tmk_phy_config> cfg;
    std::vector<uint64_t> P;         // preamble samples per link
    std::vector<uint64_t> tail_off;  // retained samples of link i: d_work[cur] + tail_off[i] .. + tail_len[i]
    std::vector<uint64_t> tail_len;
    std::vector<uint64_t> drained;   // samples removed from the front of the link's buffer so far
    std::vector<TmkLinkState> state; // buffer_offset (relative to the retained buffer), Decoding(frame_start)
    DevBuf d_work[2], d_new, d_desc;
    void* h_desc = nullptr;  // pinned staging of the copy descriptors
    size_t h_desc_cap = 0;
    int cur = 0;
    uint64_t calls = 0;
};

namespace {

struct CopyDesc {  // dst[0 .. a_len) = a[0 .. a_len), dst[a_len .. a_len + b_len) = b[0 .. b_len)
    uint64_t dst, a, a_len, b, b_len;
};

// four CTAs per link, each a quarter of the two contiguous copies; the CTAs behind them zero `pad_len` samples at `pad_at`
// (the tensor-core scan loads a tile with one TMA box that reaches 8640 samples past the tile's first sample: with the
// working buffer padded, the boxes of the last links stay inside it and no tile takes the scan's slow manual staging)
constexpr uint32_t kAsmParts = 4, kAsmPadCtas = 8;
__global__ void __launch_bounds__(256)
assemble_kernel(const CopyDesc* __restrict__ desc, uint32_t n, const float* __restrict__ src_a,
                const float* __restrict__ src_b, float* __restrict__ dst, uint64_t pad_at, uint32_t pad_len) {
    const uint32_t link = blockIdx.x / kAsmParts, part = blockIdx.x % kAsmParts;
    if (link >= n) {
        for (uint32_t k = (blockIdx.x - n * kAsmParts) * 256 + threadIdx.x; k < pad_len; k += kAsmPadCtas * 256) dst[pad_at + k] = 0.0f;
        return;
    }
    const CopyDesc d = desc[link];
    float* o = dst + d.dst;
    const float* a = src_a + d.a;
    const uint32_t k0 = part * 256 + threadIdx.x, step = kAsmParts * 256;
    for (uint64_t k = k0; k < d.a_len; k += step) o[k] = a[k];
    if (d.b_len) {
        const float* b = src_b + d.b;
        o += d.a_len;
        for (uint64_t k = k0; k < d.b_len; k += step) o[k] = b[k];
    }
}

int links_fail(tmk_links* l, int code, const char* msg) { return tmk_fail(l ? l->e : nullptr, code, msg); }

}  // namespace

extern "C" {

int tmk_links_create(tmk_engine* e, const tmk_phy_config* cfgs, uint32_t n_links, tmk_links** out) {
    if (!e || !out || (!cfgs && n_links)) return tmk_fail(e, TMK_E_INVALID, "tmk_links_create: null argument");
    tmk_links* l = new tmk_links();
    l->e = e;
    l->n = n_links;
    l->cfg.assign(cfgs, cfgs + n_links);
    l->P.resize(n_links);
    for (uint32_t i = 0; i < n_links; ++i) {
        l->cfg[i].chunk = 0;  // the feeds ARE the process_samples calls
        l->P[i] = tmk_preamble_len(&l->cfg[i]);
        if (l->P[i] == 0 || l->cfg[i].spl == 0) {
            delete l;
            return tmk_fail(e, TMK_E_INVALID, "tmk_links_create: invalid tmk_phy_config");
        }
    }
    l->tail_off.assign(n_links, 0);
    l->tail_len.assign(n_links, 0);
    l->drained.assign(n_links, 0);
    TmkLinkState s0;
    s0.off = 0;
    s0.fs = TMK_NO_FRAME;
    s0.corr = 0.f;
    s0.pad_ = 0;
    l->state.assign(n_links, s0);
    *out = l;
    return TMK_OK;
}

void tmk_links_destroy(tmk_links* l) {
    if (!l) return;
    if (l->e) {
        cudaSetDevice(l->e->device);
        cudaStreamSynchronize(l->e->stream);
    }
    l->d_work[0].release();
    l->d_work[1].release();
    l->d_new.release();
    l->d_desc.release();
    if (l->h_desc) cudaFreeHost(l->h_desc);
    delete l;
}

// PhyDecoder::reset (decoder.rs:114-119) of one link, or of all of them (link == UINT32_MAX); positions
// restart at 0 like a fresh decoder
int tmk_links_reset(tmk_links* l, uint32_t link) {
    if (!l) return TMK_E_INVALID;
    if (link != 0xffffffffu && link >= l->n) return links_fail(l, TMK_E_INVALID, "tmk_links_reset: no such link");
    for (uint32_t i = 0; i < l->n; ++i) {
        if (link != 0xffffffffu && i != link) continue;
        l->tail_len[i] = 0;
        l->drained[i] = 0;
        l->state[i].off = 0;
        l->state[i].fs = TMK_NO_FRAME;
        l->state[i].corr = 0.f;
    }
    return TMK_OK;
}

// samples buffered for link `link` (sample_buffer.len()), its buffer_offset and whether it is in Decoding state
int tmk_links_state(const tmk_links* l, uint32_t link, uint64_t* buffered, uint64_t* buffer_offset, int* decoding) {
    if (!l || link >= l->n) return TMK_E_INVALID;
    if (buffered) *buffered = l->tail_len[link];
    if (buffer_offset) *buffer_offset = l->state[link].off;
    if (decoding) *decoding = l->state[link].fs != TMK_NO_FRAME;
    return TMK_OK;
}

// PhyDecoder::process_samples (decoder.rs:68-112) for every link at once. Link i receives
// samples[offsets[i] .. offsets[i] + lens[i]) (host or device memory; a length of 0 is a call with an empty
// slice). Returns the records of the frames whose decode finished in this call, sorted by (link, order of
// detection); `stream` is the link index, positions count from the first sample ever fed to the link.
int tmk_links_process_samples(tmk_links* l, const float* samples, uint64_t total_samples, const uint64_t* offsets,
                              const uint64_t* lens, tmk_frame_rec* recs, uint64_t rec_cap, uint64_t* n_recs,
                              uint8_t* payload, uint64_t payload_cap, uint64_t* payload_len) {
    if (!l || !l->e) return TMK_E_INVALID;
    tmk_engine* e = l->e;
    if (!n_recs || !payload_len || (!recs && rec_cap) || (!payload && payload_cap) || ((!offsets || !lens) && l->n) ||
        (!samples && total_samples))
        return tmk_fail(e, TMK_E_INVALID, "tmk_links_process_samples: null argument");
    if (rec_cap && tmk_is_device_ptr(recs))
        return tmk_fail(e, TMK_E_INVALID, "tmk_links_process_samples: records are returned in host memory");
    *n_recs = 0;
    *payload_len = 0;
    const uint32_t n = l->n;
    if (n == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    for (uint32_t i = 0; i < n; ++i)
        if (offsets[i] > total_samples || lens[i] > total_samples - offsets[i])
            return tmk_fail(e, TMK_E_INVALID, "tmk_links_process_samples: chunk outside the sample buffer");

    // 1. the new samples on the device
    const float* d_new = samples;
    if (total_samples && !tmk_is_device_ptr(samples)) {
        CK(l->d_new.reserve(total_samples * sizeof(float)));
        CK(cudaMemcpyAsync(l->d_new.p, samples, total_samples * sizeof(float), cudaMemcpyHostToDevice, e->stream));
        d_new = static_cast<const float*>(l->d_new.p);
    }
    // 2. working buffers: retained tail ++ new samples, each at a multiple of 4 samples
    std::vector<CopyDesc> desc(n);
    std::vector<tmk_stream> streams(n);
    uint64_t total_work = 0;
    for (uint32_t i = 0; i < n; ++i) {
        CopyDesc& d = desc[i];
        d.dst = total_work;
        d.a = l->tail_off[i];
        d.a_len = l->tail_len[i];
        d.b = offsets[i];
        d.b_len = lens[i];
        streams[i].offset = total_work;
        streams[i].n_samples = d.a_len + d.b_len;
        streams[i].cfg = l->cfg[i];
        total_work = (total_work + streams[i].n_samples + 3) & ~3ull;
    }
    constexpr uint32_t kPad = 16384;       // zeroed samples behind the last link (see assemble_kernel)
    DevBuf& work = l->d_work[1 - l->cur];  // (the retained tails live in d_work[cur])
    CK(work.reserve((total_work + kPad + 4) * sizeof(float)));
    CK(l->d_desc.reserve(sizeof(CopyDesc) * n));
    {
        // (through pinned memory: the runtime stages a copy out of pageable memory itself, which costs more than this memcpy;
        //  the decode below ends with a synchronisation of the stream, so the buffer is free again at the next call)
        const size_t db = sizeof(CopyDesc) * n;
        const void* src = desc.data();
        if (l->h_desc_cap < db) {
            if (l->h_desc) cudaFreeHost(l->h_desc);
            l->h_desc = nullptr;
            l->h_desc_cap = 0;
            if (cudaHostAlloc(&l->h_desc, db + db / 2 + 4096, cudaHostAllocDefault) == cudaSuccess) l->h_desc_cap = db + db / 2 + 4096;
            else cudaGetLastError();
        }
        if (l->h_desc_cap >= db) {
            std::memcpy(l->h_desc, desc.data(), db);
            src = l->h_desc;
        }
        CK(cudaMemcpyAsync(l->d_desc.p, src, db, cudaMemcpyHostToDevice, e->stream));
    }
    assemble_kernel<<<n * kAsmParts + kAsmPadCtas, 256, 0, e->stream>>>(static_cast<const CopyDesc*>(l->d_desc.p), n,
                                                                        static_cast<const float*>(l->d_work[l->cur].p), d_new,
                                                                        static_cast<float*>(work.p), total_work, kPad);
    CK(cudaGetLastError());
    e->launches += 1;
    // (no synchronisation: the descriptors were copied out of pageable memory, which cudaMemcpyAsync stages before it
    //  returns, and the decode enqueues on the same stream)

    // 3. decode with the carried state
    std::vector<TmkLinkState> fin(n);
    e->link_init = l->state.data();
    e->link_final = fin.data();
    uint64_t nr = 0, np = 0;
    const int rc = tmk_decode_batch(e, static_cast<const float*>(work.p), total_work + kPad, streams.data(), n, recs, rec_cap,
                                    &nr, payload, payload_cap, &np);
    e->link_init = nullptr;
    e->link_final = nullptr;
    if (rc != TMK_OK) {  // nothing was consumed: the caller may repeat the call (e.g. with larger buffers)
        *n_recs = nr;
        *payload_len = np;
        return rc;
    }
    // 4. records: a frame still being gathered is state, not a record; positions from the link's first sample
    uint64_t w = 0;
    for (uint64_t r = 0; r < nr; ++r) {
        tmk_frame_rec rec = recs[r];
        if (rec.status == TMK_FRAME_TRUNCATED) continue;
        rec.preamble_start += l->drained[rec.stream];
        rec.frame_start += l->drained[rec.stream];
        recs[w++] = rec;
    }
    *n_recs = w;
    *payload_len = np;
    // 5. drain rule (decoder.rs:90-109): keep P-1 samples before buffer_offset — in place, in the working buffer
    for (uint32_t i = 0; i < n; ++i) {
        const uint64_t work_len = streams[i].n_samples;
        const uint64_t off = std::min<uint64_t>(fin[i].off, work_len);
        const uint64_t keep = l->P[i] - 1;
        const uint64_t drain_end = off > keep ? off - keep : 0;
        l->tail_off[i] = streams[i].offset + drain_end;
        l->tail_len[i] = work_len - drain_end;
        l->drained[i] += drain_end;
        l->state[i].off = off - drain_end;
        l->state[i].fs = fin[i].fs == TMK_NO_FRAME ? TMK_NO_FRAME : (fin[i].fs > drain_end ? fin[i].fs - drain_end : 0);
        l->state[i].corr = fin[i].corr;
    }
    l->cur = 1 - l->cur;
    l->calls++;
    return TMK_OK;
}

}  // extern "C"
