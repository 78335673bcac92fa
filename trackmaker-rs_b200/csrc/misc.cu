// misc.cu — small kernels around the hot path: modulator (PhyEncoder), correlation probe,
// channel-busy detector.
#include <algorithm>

#include "tmk_internal.cuh"

namespace {

// FOURB_FIVEB_ENCODE_TABLE, line_coding.rs:138-155
__constant__ uint8_t c_enc5[16] = {0x1e, 0x09, 0x14, 0x15, 0x0a, 0x0b, 0x0e, 0x0f,
                                   0x12, 0x13, 0x16, 0x17, 0x1a, 0x1b, 0x1c, 0x1d};

// ---------------------------------------------------------------- modulator
// One CTA per frame (PhyEncoder::encode_frame, encoder.rs:43-63): Frame::to_bytes
// (frame.rs:74-102) with CRC-8 over the payload, MSB-first bits, line code, preamble in front.
// frame_off[f] = first output sample of frame f; the inter-frame gap (encoder.rs:81-83) is
// zero-filled by the CTA of the frame in front of it (frame_off[n_frames] = total length).
constexpr int kModThreads = 256;

__device__ __forceinline__ uint32_t frame_byte(const tmk_frame_desc& fd, const uint8_t* __restrict__ payload,
                                               uint32_t crc, uint32_t i) {
    switch (i) {
        case 0: return (fd.len >> 8) & 0xffu;
        case 1: return fd.len & 0xffu;
        case 2: return crc;
        case 3: return fd.type;
        case 4: return fd.seq;
        case 5: return fd.src;
        case 6: return fd.dst;
        default: return payload[fd.payload_off + (i - 7)];
    }
}

__global__ void __launch_bounds__(kModThreads)
modulate_kernel(const TmkPattern* __restrict__ pat, const tmk_frame_desc* __restrict__ frames,
                const uint64_t* __restrict__ frame_off, const uint8_t* __restrict__ payload, float* __restrict__ out) {
    __shared__ uint32_t s_crc;
    __shared__ uint32_t s_par[kModThreads];  // NRZI level parity before each thread's span of symbols
    const uint32_t f = blockIdx.x;
    const tmk_frame_desc fd = frames[f];
    const uint32_t P = pat->P, spl = pat->spl;
    float* __restrict__ o = out + frame_off[f];
    const uint64_t span = frame_off[f + 1] - frame_off[f];
    const uint32_t nbytes = TMK_PHY_HEADER_BYTES + fd.len;

    // CRC-8 of the payload (crc.rs:7-22): sequential per thread chunk is not needed for <= 64 KiB
    // frames; one warp does it bytewise, the rest of the CTA writes the preamble meanwhile.
    if (threadIdx.x == 0) {
        uint32_t crc = 0;
        for (uint32_t i = 0; i < fd.len; ++i) crc = crc8_step(crc, payload[fd.payload_off + i]);
        s_crc = crc;
    }
    for (uint32_t i = threadIdx.x; i < P; i += kModThreads) o[i] = pat->levels[i / spl];
    __syncthreads();
    const uint32_t crc = s_crc;
    uint64_t body;
    if (pat->line_code == TMK_LINE_MANCHESTER) {
        // 0 -> [+1]*spl,[-1]*spl ; 1 -> [-1]*spl,[+1]*spl   (line_coding.rs:74-91)
        body = uint64_t(nbytes) * 16 * spl;
        for (uint64_t i = threadIdx.x; i < body; i += kModThreads) {
            const uint64_t lvl = i / spl;
            const uint32_t bit_i = uint32_t(lvl >> 1), half = uint32_t(lvl & 1);
            const uint32_t byte = frame_byte(fd, payload, crc, bit_i >> 3);
            const uint32_t bit = (byte >> (7 - (bit_i & 7))) & 1u;
            o[P + i] = (bit ^ half) ? -1.0f : 1.0f;
        }
    } else {
        // 4B5B + NRZI from level +1 (line_coding.rs:202-240). Level after symbol-bit j =
        // +1 * (-1)^(number of 1 bits so far): prefix parity over the 5-bit symbols.
        const uint32_t nsym = nbytes * 2;
        const uint32_t per = (nsym + kModThreads - 1) / kModThreads;
        const uint32_t s0 = threadIdx.x * per, s1 = min(nsym, s0 + per);
        uint32_t par = 0;
        for (uint32_t s = s0; s < s1; ++s) {
            const uint32_t byte = frame_byte(fd, payload, crc, s >> 1);
            const uint32_t nib = (s & 1) ? (byte & 0xf) : (byte >> 4);
            par ^= __popc(c_enc5[nib]) & 1u;
        }
        s_par[threadIdx.x] = par;
        __syncthreads();
        if (threadIdx.x == 0) {  // exclusive prefix parity over the thread spans
            uint32_t acc = 0;
            for (int t = 0; t < kModThreads; ++t) {
                const uint32_t v = s_par[t];
                s_par[t] = acc;
                acc ^= v;
            }
        }
        __syncthreads();
        par = s_par[threadIdx.x];
        body = uint64_t(nsym) * 5 * spl;
        for (uint32_t s = s0; s < s1; ++s) {
            const uint32_t byte = frame_byte(fd, payload, crc, s >> 1);
            const uint32_t sym = c_enc5[(s & 1) ? (byte & 0xf) : (byte >> 4)];
            for (int j = 4; j >= 0; --j) {
                par ^= (sym >> j) & 1u;
                const float level = par ? -1.0f : 1.0f;
                float* dst = o + P + (uint64_t(s) * 5 + (4 - j)) * spl;
                for (uint32_t k = 0; k < spl; ++k) dst[k] = level;
            }
        }
    }
    for (uint64_t i = P + body + threadIdx.x; i < span; i += kModThreads) o[i] = 0.0f;  // gap
}

// ---------------------------------------------------------------- correlation probe
// One search_for_preamble pass that never stops (decoder.rs:130-147,224-233): window energy seeded
// at window 0 by a sequential sum and ROLLED from there, AVX-order dot product. Bit-exact with the
// reference arithmetic, so it is sequential per stream: kernel 1 rolls the energy chain (one warp
// per stream, 32 steps per coalesced load), kernel 2 divides the dot products by it in place.
__global__ void energy_chain_kernel(const float* __restrict__ samples, const TmkStreamDev* __restrict__ streams,
                                    uint32_t n_streams, const TmkPattern* __restrict__ patterns,
                                    float* __restrict__ corr_out) {
    const uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (s >= n_streams) return;
    const TmkStreamDev sd = streams[s];
    const uint32_t P = patterns[sd.pattern].P;
    if (sd.n_samples < P) return;
    const uint64_t nwin = sd.n_samples - P + 1;
    const float* __restrict__ x = samples + sd.offset;
    float* __restrict__ out = corr_out + sd.offset;
    float e = 0.0f;
    for (uint32_t i = 0; i < P; ++i) e = ex_add(e, ex_mul(x[i], x[i]));
    for (uint64_t w0 = 0; w0 < nwin; w0 += 32) {
        const uint64_t w = w0 + lane;
        const bool step_ok = w + 1 < nwin;  // the roll that produces window w + 1
        const float lv = step_ok ? x[w] : 0.f, en = step_ok ? x[w + P] : 0.f;
        float mine = 0.f;
        for (int j = 0; j < 32; ++j) {
            if (lane == j) mine = e;  // energy of window w0 + j
            if (w0 + j + 1 < nwin) e = ex_roll(e, __shfl_sync(0xffffffffu, lv, j), __shfl_sync(0xffffffffu, en, j));
        }
        if (w < nwin) out[w] = mine;
    }
}

__global__ void correlate_kernel(const float* __restrict__ samples, const TmkStreamDev* __restrict__ streams,
                                 const TmkPattern* __restrict__ patterns, float* __restrict__ corr_out) {
    const TmkStreamDev sd = streams[blockIdx.x];
    const TmkPattern* pat = patterns + sd.pattern;
    const uint32_t P = pat->P, spl = pat->spl;
    if (sd.n_samples < P) return;
    const uint64_t nwin = sd.n_samples - P + 1;
    const float* __restrict__ x = samples + sd.offset;
    for (uint64_t w = uint64_t(blockIdx.y) * blockDim.x + threadIdx.x; w < nwin; w += uint64_t(gridDim.y) * blockDim.x) {
        float a[8];
#pragma unroll
        for (int l = 0; l < 8; ++l) a[l] = 0.0f;
        uint32_t i = 0;
        for (; i + 8 <= P; i += 8) {
#pragma unroll
            for (int l = 0; l < 8; ++l) a[l] = ex_add(a[l], ex_mul(x[w + i + l], pat->levels[(i + l) / spl]));
        }
        float s = 0.0f;
#pragma unroll
        for (int l = 0; l < 4; ++l) s = ex_add(s, ex_add(a[l], a[l + 4]));
        for (; i < P; ++i) s = ex_add(s, ex_mul(x[w + i], pat->levels[i / spl]));
        corr_out[sd.offset + w] = ex_corr(s, corr_out[sd.offset + w], pat->pe);
    }
}

// ---------------------------------------------------------------- mac::is_channel_busy (src/mac/mod.rs:18-27)
__global__ void busy_kernel(const float* __restrict__ samples, const TmkStreamDev* __restrict__ streams,
                            int8_t* __restrict__ busy) {
    const TmkStreamDev sd = streams[blockIdx.x];
    if (sd.n_samples < 20) {
        if (threadIdx.x == 0) busy[blockIdx.x] = -1;
        return;
    }
    const float* __restrict__ x = samples + sd.offset;
    int any = 0;
    for (uint64_t i = threadIdx.x; i < sd.n_samples && !any; i += blockDim.x) any |= fabsf(x[i]) > 0.5f;
    any = __syncthreads_or(any);
    if (threadIdx.x == 0) busy[blockIdx.x] = any ? 1 : 0;
}

// ---------------------------------------------------------------- 16-bit PCM
// convert_audio_buffer_to_f32, S16 arm (src/audio/codec.rs:121-125): sample as f32 / 32768.0 — a power-of-two scale,
// exact. Four samples per thread per step: one 8-byte load, one 16-byte store, grid-stride over 148 x 8 CTAs.
__global__ void __launch_bounds__(256) pcm16_to_f32_kernel(const int16_t* __restrict__ pcm, float* __restrict__ out, uint64_t n) {
    const uint64_t n4 = n / 4, stride = uint64_t(gridDim.x) * blockDim.x;
    const short4* __restrict__ in4 = reinterpret_cast<const short4*>(pcm);
    float4* __restrict__ out4 = reinterpret_cast<float4*>(out);
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const short4 v = __ldg(in4 + i);
        out4[i] = make_float4(float(v.x) * (1.0f / 32768.0f), float(v.y) * (1.0f / 32768.0f), float(v.z) * (1.0f / 32768.0f),
                              float(v.w) * (1.0f / 32768.0f));
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) out[n4 * 4 + threadIdx.x] = float(pcm[n4 * 4 + threadIdx.x]) * (1.0f / 32768.0f);
}

// dump_to_wav (src/utils/dump.rs:37-40): (sample * i16::MAX as f32) as i16. Rust's float -> int `as` truncates toward
// zero, saturates, and maps NaN to 0: cvt.rzi.s32.f32 does the same at 32 bits, then the clamp to 16.
__device__ __forceinline__ int16_t pcm16_of(float s) {
    const int v = __float2int_rz(__fmul_rn(s, 32767.0f));
    return int16_t(max(-32768, min(32767, v)));
}

__global__ void __launch_bounds__(256) pcm16_encode_kernel(const float* __restrict__ x, int16_t* __restrict__ out, uint64_t n) {
    const uint64_t n4 = n / 4, stride = uint64_t(gridDim.x) * blockDim.x;
    const float4* __restrict__ in4 = reinterpret_cast<const float4*>(x);
    short4* __restrict__ out4 = reinterpret_cast<short4*>(out);
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 v = __ldg(in4 + i);
        out4[i] = make_short4(pcm16_of(v.x), pcm16_of(v.y), pcm16_of(v.z), pcm16_of(v.w));
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) out[n4 * 4 + threadIdx.x] = pcm16_of(x[n4 * 4 + threadIdx.x]);
}

}  // namespace

int tmk_launch_pcm16_to_f32(const int16_t* pcm, float* out, uint64_t n, cudaStream_t st) {
    if (n == 0) return 0;
    const uint64_t want = (n / 4 + 255) / 256;
    pcm16_to_f32_kernel<<<unsigned(std::min<uint64_t>(std::max<uint64_t>(want, 1), 148 * 8)), 256, 0, st>>>(pcm, out, n);
    return 1;
}

int tmk_launch_pcm16_encode(const float* samples, int16_t* out, uint64_t n, cudaStream_t st) {
    if (n == 0) return 0;
    const uint64_t want = (n / 4 + 255) / 256;
    pcm16_encode_kernel<<<unsigned(std::min<uint64_t>(std::max<uint64_t>(want, 1), 148 * 8)), 256, 0, st>>>(samples, out, n);
    return 1;
}

int tmk_launch_modulate(const TmkPattern* pattern, const tmk_frame_desc* frames, const uint64_t* frame_off,
                        uint32_t n_frames, const uint8_t* payload, float* out, cudaStream_t st) {
    if (n_frames == 0) return 0;
    modulate_kernel<<<n_frames, kModThreads, 0, st>>>(pattern, frames, frame_off, payload, out);
    return 1;
}

int tmk_launch_correlate(const float* samples, const TmkStreamDev* streams, uint32_t n_streams,
                         const TmkPattern* patterns, float* corr_out, cudaStream_t st) {
    if (n_streams == 0) return 0;
    energy_chain_kernel<<<(n_streams + 3) / 4, 128, 0, st>>>(samples, streams, n_streams, patterns, corr_out);
    dim3 grid(n_streams, 64);
    correlate_kernel<<<grid, 256, 0, st>>>(samples, streams, patterns, corr_out);
    return 2;
}

int tmk_launch_busy(const float* samples, const TmkStreamDev* streams, uint32_t n_streams, int8_t* busy,
                    cudaStream_t st) {
    if (n_streams == 0) return 0;
    busy_kernel<<<n_streams, 256, 0, st>>>(samples, streams, busy);
    return 1;
}
