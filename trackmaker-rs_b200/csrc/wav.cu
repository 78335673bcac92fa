// wav.cu — capture container ingest: RIFF/WAVE files as the reference stores its recordings
// (dump_to_wav, src/utils/dump.rs:23-46: hound 3.5.1 writes a `fmt ` chunk — 16 bytes of PCM header, or the
// 40-byte WAVE_FORMAT_EXTENSIBLE form — and one `data` chunk) turned into the f32 samples its reader produces
// (convert_audio_buffer_to_f32, src/audio/codec.rs:114-148: CHANNEL 0 only; U8 (s - 128) / 128, S16 s / 32768,
// S24 s / 8388608, S32 s / 2147483648, F32 as is — all scalings are powers of two, exact in f32).
// The chunk walk is host work (a few dozen bytes per file); the samples cross PCIe as the raw file bytes and are
// unpacked by a kernel in front of the decode.
#include <cstring>
#include <string>
#include <vector>

#include "engine.hpp"

namespace {

struct WavDesc {
    uint64_t src;       // byte offset of the first frame in the blob
    uint64_t dst;       // first output sample
    uint64_t n_frames;  // samples per channel
    uint32_t stride;    // bytes per frame (block align)
    uint32_t kind;      // 0 U8, 1 S16, 2 S24, 3 S32, 4 F32
};

__device__ __forceinline__ float wav_sample(const uint8_t* __restrict__ p, uint32_t kind) {
    switch (kind) {
        case 0: return (float(p[0]) - 128.0f) / 128.0f;
        case 1: return float(int16_t(uint16_t(p[0]) | (uint16_t(p[1]) << 8))) / 32768.0f;
        case 2: {
            const int32_t v = int32_t((uint32_t(p[0]) << 8) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 24)) >> 8;
            return float(v) / 8388608.0f;
        }
        case 3: {
            const int32_t v = int32_t(uint32_t(p[0]) | (uint32_t(p[1]) << 8) | (uint32_t(p[2]) << 16) | (uint32_t(p[3]) << 24));
            return float(v) / 2147483648.0f;  // `sample as f32`: round to nearest, then an exact scaling
        }
        default: return __uint_as_float(uint32_t(p[0]) | (uint32_t(p[1]) << 8) | (uint32_t(p[2]) << 16) | (uint32_t(p[3]) << 24));
    }
}

// one CTA row per file (grid.y), grid-stride over its frames
__global__ void __launch_bounds__(256)
wav_unpack_kernel(const uint8_t* __restrict__ blob, const WavDesc* __restrict__ desc, float* __restrict__ out) {
    const WavDesc d = desc[blockIdx.y];
    const uint8_t* __restrict__ src = blob + d.src;
    float* __restrict__ o = out + d.dst;
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < d.n_frames; i += uint64_t(gridDim.x) * blockDim.x)
        o[i] = wav_sample(src + i * d.stride, d.kind);
}

uint32_t rd32(const uint8_t* p) { return uint32_t(p[0]) | (uint32_t(p[1]) << 8) | (uint32_t(p[2]) << 16) | (uint32_t(p[3]) << 24); }
uint32_t rd16(const uint8_t* p) { return uint32_t(p[0]) | (uint32_t(p[1]) << 8); }

}  // namespace

extern "C" {

// RIFF chunk walk. Returns TMK_OK or TMK_E_INVALID (not a WAVE file, no `fmt ` / `data` chunk, unsupported sample type).
int tmk_wav_parse(const uint8_t* file, uint64_t len, tmk_wav_info* out) {
    if (!file || !out) return TMK_E_INVALID;
    std::memset(out, 0, sizeof *out);
    if (len < 12 || std::memcmp(file, "RIFF", 4) != 0 || std::memcmp(file + 8, "WAVE", 4) != 0) return TMK_E_INVALID;
    bool have_fmt = false;
    uint64_t pos = 12;
    while (pos + 8 <= len) {
        const uint8_t* ck = file + pos;
        const uint64_t sz = rd32(ck + 4);
        const uint64_t body = pos + 8;
        if (std::memcmp(ck, "fmt ", 4) == 0) {
            if (sz < 16 || body + 16 > len) return TMK_E_INVALID;
            uint32_t fmt = rd16(file + body);
            out->channels = rd16(file + body + 2);
            out->sample_rate = rd32(file + body + 4);
            out->block_align = rd16(file + body + 12);
            out->bits_per_sample = rd16(file + body + 14);
            if (fmt == 0xFFFEu) {  // WAVE_FORMAT_EXTENSIBLE: the sub-format GUID starts with the format tag
                if (sz < 40 || body + 40 > len) return TMK_E_INVALID;
                fmt = rd16(file + body + 24);
            }
            out->format = fmt;
            have_fmt = true;
        } else if (std::memcmp(ck, "data", 4) == 0) {
            if (!have_fmt) return TMK_E_INVALID;
            uint64_t n = sz;
            if (body + n > len) n = len - body;  // a writer that was not finalised leaves a short file
            if (out->block_align == 0 || out->channels == 0) return TMK_E_INVALID;
            out->data_off = body;
            out->n_frames = n / out->block_align;
            const uint32_t bps = out->bits_per_sample;
            const bool ok = (out->format == 1 && (bps == 8 || bps == 16 || bps == 24 || bps == 32)) || (out->format == 3 && bps == 32);
            if (!ok || out->block_align < out->channels * (bps / 8)) return TMK_E_INVALID;
            return TMK_OK;
        }
        pos = body + sz + (sz & 1);  // chunks are word aligned
    }
    return TMK_E_INVALID;
}

}  // extern "C"

namespace {

// parse every file of the blob: descriptors, per-file sample offsets (multiples of 4) and lengths, total samples
int wav_layout(tmk_engine* e, const uint8_t* blob, uint64_t blob_len, const uint64_t* file_off, uint32_t n_files,
               std::vector<WavDesc>* desc, uint64_t* total_out) {
    const bool dev_blob = blob_len && tmk_is_device_ptr(blob);
    desc->resize(n_files);
    std::vector<uint8_t> head;
    uint64_t total = 0;
    for (uint32_t i = 0; i < n_files; ++i) {
        if (file_off[i] > file_off[i + 1] || file_off[i + 1] > blob_len)
            return tmk_fail(e, TMK_E_INVALID, "wav: file offsets out of range");
        const uint64_t flen = file_off[i + 1] - file_off[i];
        tmk_wav_info info;
        int rc;
        if (dev_blob) {  // of a device blob only the first 4 KiB of every file are fetched for the chunk walk
            const uint64_t hl = flen < 4096 ? flen : 4096;
            head.assign(hl, 0);
            CK(cudaMemcpy(head.data(), blob + file_off[i], hl, cudaMemcpyDeviceToHost));
            rc = tmk_wav_parse(head.data(), hl, &info);
            if (rc == TMK_OK) {  // the data chunk runs past the fetched head: its length comes from the file size
                const uint64_t avail = flen - info.data_off;
                const uint64_t want = uint64_t(rd32(head.data() + info.data_off - 4));
                info.n_frames = (want < avail ? want : avail) / info.block_align;
            }
        } else {
            rc = tmk_wav_parse(blob + file_off[i], flen, &info);
        }
        if (rc != TMK_OK) return tmk_fail(e, TMK_E_INVALID, "wav: file " + std::to_string(i) + " is not a supported WAVE file");
        WavDesc& d = (*desc)[i];
        d.src = file_off[i] + info.data_off;
        d.dst = total;
        d.n_frames = info.n_frames;
        d.stride = info.block_align;
        d.kind = info.format == 3 ? 4u : info.bits_per_sample == 8 ? 0u : info.bits_per_sample == 16 ? 1u : info.bits_per_sample == 24 ? 2u : 3u;
        total = (total + info.n_frames + 3) & ~3ull;
    }
    *total_out = total;
    return TMK_OK;
}

// unpack channel 0 of every file into d_out (device memory, `total` floats)
int wav_unpack(tmk_engine* e, const uint8_t* blob, uint64_t blob_len, const std::vector<WavDesc>& desc, uint64_t total, float* d_out) {
    const uint32_t n_files = uint32_t(desc.size());
    const uint8_t* d_blob = blob;
    if (!tmk_is_device_ptr(blob)) {
        CK(e->d_wav.reserve(blob_len));
        CK(cudaMemcpyAsync(e->d_wav.p, blob, blob_len, cudaMemcpyHostToDevice, e->stream));
        d_blob = static_cast<const uint8_t*>(e->d_wav.p);
    }
    CK(e->d_wavdesc.reserve(sizeof(WavDesc) * n_files));
    CK(cudaMemcpyAsync(e->d_wavdesc.p, desc.data(), sizeof(WavDesc) * n_files, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemsetAsync(d_out, 0, total * sizeof(float), e->stream));  // the alignment padding between streams
    uint64_t longest = 0;
    for (const WavDesc& d : desc) longest = d.n_frames > longest ? d.n_frames : longest;
    uint32_t gx = uint32_t(std::min<uint64_t>((longest + 255) / 256, 148ull * 8));
    if (gx == 0) gx = 1;
    for (uint32_t f0 = 0; f0 < n_files; f0 += 65535) {  // grid.y limit
        const uint32_t nf = std::min<uint32_t>(65535u, n_files - f0);
        wav_unpack_kernel<<<dim3(gx, nf), 256, 0, e->stream>>>(d_blob, static_cast<const WavDesc*>(e->d_wavdesc.p) + f0, d_out);
        e->launches += 1;
    }
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(e->stream));  // `desc` is the caller's
    return TMK_OK;
}

}  // namespace

extern "C" {

// files: n_files WAVE files back to back, file i = blob[file_off[i] .. file_off[i + 1]) (blob: host or device; file_off:
// host, n_files + 1 entries). Channel 0 of file i goes to samples_out[offsets_out[i] .. + lens_out[i]) (offsets are
// multiples of 4 samples, packed in file order). samples_out: device or host memory of sample_cap floats; *total_out
// is always the size needed (call with sample_cap = 0 to size the buffer).
int tmk_wav_to_f32(tmk_engine* e, const uint8_t* blob, uint64_t blob_len, const uint64_t* file_off, uint32_t n_files,
                   float* samples_out, uint64_t sample_cap, uint64_t* offsets_out, uint64_t* lens_out, uint64_t* total_out) {
    if (!e || (!blob && blob_len) || (!file_off && n_files) || !total_out)
        return tmk_fail(e, TMK_E_INVALID, "tmk_wav_to_f32: null argument");
    CK(cudaSetDevice(e->device));
    std::vector<WavDesc> desc;
    uint64_t total = 0;
    int rc = wav_layout(e, blob, blob_len, file_off, n_files, &desc, &total);
    if (rc != TMK_OK) return rc;
    for (uint32_t i = 0; i < n_files; ++i) {
        if (offsets_out) offsets_out[i] = desc[i].dst;
        if (lens_out) lens_out[i] = desc[i].n_frames;
    }
    *total_out = total;
    if (total > sample_cap) return tmk_fail(e, TMK_E_CAPACITY, "tmk_wav_to_f32: sample buffer too small");
    if (n_files == 0 || total == 0) return TMK_OK;
    if (!samples_out) return tmk_fail(e, TMK_E_INVALID, "tmk_wav_to_f32: null output");
    float* d_out = samples_out;
    const bool dev_out = tmk_is_device_ptr(samples_out);
    if (!dev_out) {
        CK(e->d_samples.reserve(total * sizeof(float)));
        d_out = static_cast<float*>(e->d_samples.p);
    }
    rc = wav_unpack(e, blob, blob_len, desc, total, d_out);
    if (rc != TMK_OK) return rc;
    if (!dev_out) {
        CK(cudaMemcpyAsync(samples_out, d_out, total * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
    }
    return TMK_OK;
}

// WAVE files in, verdict records out: the unpack kernel in front of tmk_decode_batch. cfgs: one per file.
// streams_out (optional, n_files entries): the stream table of the batch (offset / n_samples of every file's channel 0;
// the records' `stream` is the file index).
int tmk_decode_batch_wav(tmk_engine* e, const uint8_t* blob, uint64_t blob_len, const uint64_t* file_off, uint32_t n_files,
                         const tmk_phy_config* cfgs, tmk_stream* streams_out, tmk_frame_rec* recs, uint64_t rec_cap,
                         uint64_t* n_recs, uint8_t* payload, uint64_t payload_cap, uint64_t* payload_len) {
    if (!e || !n_recs || !payload_len || (!cfgs && n_files) || (!blob && blob_len) || (!file_off && n_files))
        return tmk_fail(e, TMK_E_INVALID, "tmk_decode_batch_wav: null argument");
    *n_recs = 0;
    *payload_len = 0;
    if (n_files == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    std::vector<WavDesc> desc;
    uint64_t total = 0;
    int rc = wav_layout(e, blob, blob_len, file_off, n_files, &desc, &total);
    if (rc != TMK_OK) return rc;
    std::vector<tmk_stream> st(n_files);
    for (uint32_t i = 0; i < n_files; ++i) {
        st[i].offset = desc[i].dst;
        st[i].n_samples = desc[i].n_frames;
        st[i].cfg = cfgs[i];
    }
    if (streams_out) std::memcpy(streams_out, st.data(), sizeof(tmk_stream) * n_files);
    if (total == 0) return TMK_OK;
    CK(e->d_pcm.reserve(total * sizeof(float)));  // f32 samples of the batch (the PCM scratch is free here)
    float* d_x = static_cast<float*>(e->d_pcm.p);
    rc = wav_unpack(e, blob, blob_len, desc, total, d_x);
    if (rc != TMK_OK) return rc;
    return tmk_decode_batch(e, d_x, total, st.data(), n_files, recs, rec_cap, n_recs, payload, payload_cap, payload_len);
}

}  // extern "C"
