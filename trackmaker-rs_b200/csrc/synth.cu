// synth.cu — synthetic capture generator for the measurement harness: every stream
// is AWGN plus randomly placed src/phy frames (random payloads, 10 % ACKs, 10 % foreign dst,
// gaps U[48,4800)), generated on the GPU from counter-based hashes of (seed, stream, index).
// The frames are produced exactly as PhyEncoder would (encoder.rs:43-63): header + CRC-8 +
// line code behind the preamble, scaled by `amplitude` and added onto the noise.
// Impairments of SURVEY §8d (tmk_synth_batch_ex): 3-tap multipath h = [1, 0.35 @ t1, -0.2 @ t2] renormalised to
// unit energy (t1 in U{2..12}, t2 in U{13..40} per stream), clock drift (linear-interpolation resample by 1 + delta,
// delta in U[-ppm, +ppm] per stream), per-stream amplitude scale U[amp_min, 1], sparse frame plans (gap range).
// Also the chirp + BPSK capture generator of examples/ask.rs frames (tmk_ask_synth_batch).
#include "tmk_internal.cuh"

namespace {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// per-stream impairment parameters, derived from (seed, stream) only
struct StreamImp {
    float scale;     // amplitude scale of the whole capture (signal and noise)
    int tau1, tau2;  // echo delays in samples
    float h0, h1, h2;
    double rate;     // 1 + delta: output sample n reads the source signal at t = n * rate
};
__device__ __forceinline__ StreamImp stream_imp(uint64_t seed, uint32_t s, const TmkSynthImp& imp) {
    StreamImp r;
    const uint64_t h = mix64(seed ^ (uint64_t(s) << 24) ^ 0xD21F7ull);
    const float u0 = float(uint32_t(h) & 0xffffffu) * (1.0f / 16777216.0f);
    const float u1 = float(uint32_t(h >> 24) & 0xffffffu) * (1.0f / 16777216.0f);
    r.scale = imp.amp_min > 0.0f ? imp.amp_min + (1.0f - imp.amp_min) * u0 : 1.0f;
    r.tau1 = 2 + int((h >> 48) % 11);
    r.tau2 = 13 + int((h >> 56) % 28);
    if (imp.multipath) {
        r.h0 = 0.92747675f;   // 1 / sqrt(1 + 0.35^2 + 0.2^2)
        r.h1 = 0.32461686f;   // 0.35 / ...
        r.h2 = -0.18549535f;  // -0.2 / ...
    } else {
        r.h0 = 1.0f;
        r.h1 = r.h2 = 0.0f;
    }
    r.rate = imp.drift_ppm != 0.0f ? 1.0 + double(imp.drift_ppm) * 1e-6 * double(2.0f * u1 - 1.0f) : 1.0;
    return r;
}
// source samples a frame plan may use: the echo tail and the resampling must stay inside the capture
__device__ __forceinline__ uint64_t plan_limit(uint64_t n_per_stream, const StreamImp& si, const TmkSynthImp& imp) {
    uint64_t lim = n_per_stream;
    if (imp.drift_ppm != 0.0f) lim = uint64_t(double(n_per_stream - 2) * si.rate);
    const uint64_t tail = imp.multipath ? 41 : 0;
    return lim > tail ? lim - tail : 0;
}

__constant__ uint8_t c_enc5s[16] = {0x1e, 0x09, 0x14, 0x15, 0x0a, 0x0b, 0x0e, 0x0f,
                                    0x12, 0x13, 0x16, 0x17, 0x1a, 0x1b, 0x1c, 0x1d};

// 4 samples per thread: two Box-Muller pairs
__global__ void noise_kernel(uint64_t seed, uint64_t total, float sigma, uint64_t n_per_stream, TmkSynthImp imp,
                             float* __restrict__ out) {
    const uint64_t i4 = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
    if (i4 >= total) return;
    if (imp.amp_min > 0.0f) sigma *= stream_imp(seed, uint32_t(i4 / n_per_stream), imp).scale;
    float v[4];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const uint64_t h = mix64(seed ^ mix64(i4 / 2 + p));
        const float u1 = (float(uint32_t(h >> 40)) + 1.0f) * (1.0f / 16777216.0f);
        const float u2 = float(uint32_t(h) & 0xffffffu) * (1.0f / 16777216.0f);
        const float r = sigma * sqrtf(-2.0f * __logf(u1));
        float sn, cs;
        __sincosf(6.2831853f * u2, &sn, &cs);
        v[2 * p] = r * cs;
        v[2 * p + 1] = r * sn;
    }
    if (i4 + 3 < total) {
        *reinterpret_cast<float4*>(out + i4) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
        for (int k = 0; k < 4 && i4 + k < total; ++k) out[i4 + k] = v[k];
    }
}

// One thread per stream walks its frame plan. With frames == nullptr it only counts.
__global__ void plan_kernel(const TmkPattern* __restrict__ pat, int32_t local_addr, uint64_t seed, uint32_t n_streams,
                            uint64_t n_per_stream, TmkSynthImp imp, uint32_t* __restrict__ counts,
                            const uint64_t* __restrict__ base, tmk_frame_rec* __restrict__ frames, uint64_t frame_cap) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    const StreamImp si = stream_imp(seed, s, imp);
    const uint64_t limit = plan_limit(n_per_stream, si, imp);
    uint64_t st = mix64(seed ^ (uint64_t(s) << 20) ^ 0x5151ull);
    auto next = [&]() { st = mix64(st); return st; };
    uint64_t pos = next() % 4800;
    uint32_t k = 0, seq = 0;
    const uint8_t me = local_addr >= 0 ? uint8_t(local_addr) : 1;
    for (;;) {
        const uint64_t r = next();
        const bool ack = (r % 10) == 0;
        const uint32_t len = ack ? 0 : 1 + uint32_t((r >> 8) % 128);
        const bool foreign = ((r >> 24) % 10) == 0;
        const uint8_t dst = foreign ? uint8_t(me + 1 + ((r >> 32) % 200)) : me;
        const uint8_t src = uint8_t(r >> 48);
        const uint64_t flen = pat->P + tmk_sfb(pat->line_code, pat->spl, 8ull * (TMK_PHY_HEADER_BYTES + len));
        if (pos + flen > limit) break;
        if (frames) {
            const uint64_t slot = base[s] + k;
            if (slot < frame_cap) {
                tmk_frame_rec t;
                t.stream = s;
                t.status = TMK_FRAME_OK;
                t.preamble_start = pos;  // in SOURCE samples (frames_kernel resamples; == capture samples without drift)
                t.frame_start = pos + pat->P;
                t.payload_off = 0;
                t.corr = 1.0f;
                t.consumed = uint32_t(flen);
                t.len = uint16_t(len);
                t.crc = 0;  // filled by frames_kernel
                t.type = ack ? TMK_FRAME_TYPE_ACK : TMK_FRAME_TYPE_DATA;
                t.seq = uint8_t(seq);
                t.src = src;
                t.dst = dst;
                t.hdr_valid = 1;
                frames[slot] = t;
            }
        }
        ++k;
        ++seq;
        pos += flen + imp.gap_min + next() % imp.gap_span;
    }
    if (!frames) counts[s] = k;
}

__device__ __forceinline__ uint8_t payload_byte(uint64_t seed, uint32_t stream, uint64_t pos, uint32_t i) {
    return uint8_t(mix64(seed ^ (uint64_t(stream) << 40) ^ (pos << 8) ^ 0xABCDull ^ (uint64_t(i) << 52)) >> 17);
}

// One CTA per planned frame: build the frame bytes, then add amplitude * (the line-coded levels through the
// stream's multipath FIR, read at the drifted sample clock) onto the capture. Without impairments that is exactly
// amplitude * level per sample.
__global__ void __launch_bounds__(128)
frames_kernel(const TmkPattern* __restrict__ pat, uint64_t seed, uint64_t n_per_stream, float amplitude, TmkSynthImp imp,
              tmk_frame_rec* __restrict__ frames, float* __restrict__ samples) {
    __shared__ uint8_t fb[TMK_PHY_HEADER_BYTES + 128];
    __shared__ uint8_t lvl[(TMK_PHY_HEADER_BYTES + 128) * 16];  // level signs of the body (1 = negative)
    const tmk_frame_rec fr = frames[blockIdx.x];
    const uint32_t len = fr.len, nbytes = TMK_PHY_HEADER_BYTES + len;
    for (uint32_t i = threadIdx.x; i < len; i += blockDim.x)
        fb[TMK_PHY_HEADER_BYTES + i] = payload_byte(seed, fr.stream, fr.preamble_start, i);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t crc = 0;
        for (uint32_t i = 0; i < len; ++i) crc = crc8_step(crc, fb[TMK_PHY_HEADER_BYTES + i]);
        fb[0] = uint8_t(len >> 8);
        fb[1] = uint8_t(len);
        fb[2] = uint8_t(crc);
        fb[3] = fr.type;
        fb[4] = fr.seq;
        fb[5] = fr.src;
        fb[6] = fr.dst;
        frames[blockIdx.x].crc = uint8_t(crc);
    }
    __syncthreads();
    uint32_t nlev;
    if (pat->line_code == TMK_LINE_MANCHESTER) {
        nlev = nbytes * 16;
        for (uint32_t l = threadIdx.x; l < nlev; l += blockDim.x) {
            const uint32_t bi = l >> 1, bit = (fb[bi >> 3] >> (7 - (bi & 7))) & 1u;
            lvl[l] = uint8_t(bit ^ (l & 1));
        }
    } else {
        nlev = nbytes * 10;
        if (threadIdx.x == 0) {
            uint32_t par = 0;
            for (uint32_t s = 0; s < nbytes * 2; ++s) {
                const uint32_t sym = c_enc5s[(s & 1) ? (fb[s >> 1] & 0xf) : (fb[s >> 1] >> 4)];
                for (int j = 4; j >= 0; --j) {
                    par ^= (sym >> j) & 1u;
                    lvl[s * 5 + (4 - j)] = uint8_t(par);
                }
            }
        }
    }
    __syncthreads();
    const uint32_t P = pat->P, spl = pat->spl;
    const int64_t total = int64_t(P) + int64_t(nlev) * spl;
    const StreamImp si = stream_imp(seed, fr.stream, imp);
    const float amp = amplitude * si.scale;
    float* __restrict__ o = samples + uint64_t(fr.stream) * n_per_stream;
    const int64_t pos = int64_t(fr.preamble_start);
    auto level = [&](int64_t j) -> float {  // line-coded level of source sample pos + j (0 outside the frame)
        if (j < 0 || j >= total) return 0.0f;
        return j < int64_t(P) ? pat->levels[uint32_t(j) / spl] : (lvl[uint32_t(j - P) / spl] ? -1.0f : 1.0f);
    };
    auto clean = [&](int64_t j) -> float {  // through the multipath FIR
        float v = si.h0 * level(j);
        if (imp.multipath) v += si.h1 * level(j - si.tau1) + si.h2 * level(j - si.tau2);
        return v;
    };
    // capture samples this frame reaches: source [pos, pos + total + tail) at the drifted clock
    const int64_t tail = imp.multipath ? 41 : 0;
    int64_t n_lo = int64_t(double(pos) / si.rate) - 1, n_hi = int64_t(double(pos + total + tail) / si.rate) + 2;
    if (n_lo < 0) n_lo = 0;
    if (n_hi > int64_t(n_per_stream)) n_hi = int64_t(n_per_stream);
    for (int64_t n = n_lo + threadIdx.x; n < n_hi; n += blockDim.x) {
        const double t = double(n) * si.rate;
        const int64_t i0 = int64_t(t);
        const float frac = float(t - double(i0));
        const float v = clean(i0 - pos) * (1.0f - frac) + clean(i0 + 1 - pos) * frac;
        if (v != 0.0f) o[n] += amp * v;
    }
    if (threadIdx.x == 0 && imp.drift_ppm != 0.0f) {  // truth positions in capture samples
        frames[blockIdx.x].preamble_start = uint64_t(double(pos) / si.rate + 0.5);
        frames[blockIdx.x].frame_start = uint64_t(double(pos + P) / si.rate + 0.5);
    }
}

// ---------------------------------------------------------------- chirp + BPSK captures (examples/ask.rs:608-637)
// One CTA per stream: frames of 440 preamble samples + 108 bits x 44 samples (id = k % 100 + 1 in the first 8 bits,
// 92 hashed payload bits, 8 zero bits), separated by gaps U[0,100) + U[0,100) like the reference's TX loop, added
// onto the noise. tab = [440 preamble | 4752 carrier] (the engine's host-computed tables).
__global__ void __launch_bounds__(256)
ask_synth_kernel(const float* __restrict__ tab, uint64_t seed, uint64_t n_per_stream, float* __restrict__ samples,
                 unsigned long long* __restrict__ n_frames) {
    const uint32_t s = blockIdx.x;
    float* __restrict__ o = samples + uint64_t(s) * n_per_stream;
    const float* __restrict__ pre = tab;
    const float* __restrict__ car = tab + TMK_ASK_PREAMBLE;
    constexpr uint32_t kLen = TMK_ASK_PREAMBLE + TMK_ASK_FRAME_SAMPLES;
    uint64_t st = mix64(seed ^ (uint64_t(s) << 20) ^ 0xA5Cull);
    uint64_t pos = 0;
    uint32_t k = 0;
    for (;; ++k) {
        st = mix64(st);
        const uint64_t r = st;
        pos += r % 100;  // gap1
        if (pos + kLen + 2 > n_per_stream) break;
        const uint32_t id = k % 100 + 1;
        for (uint32_t i = threadIdx.x; i < kLen; i += blockDim.x) {
            float v;
            if (i < TMK_ASK_PREAMBLE) {
                v = pre[i];
            } else {
                const uint32_t kk = i - TMK_ASK_PREAMBLE, j = kk / TMK_ASK_SAMPLES_PER_BIT;
                bool one = false;
                if (j < 8) one = (id >> (7 - j)) & 1u;
                else if (j < TMK_ASK_DATA_BITS) one = (mix64(r ^ (uint64_t(j) << 32)) >> 13) & 1u;
                v = one ? car[kk] : -car[kk];
            }
            o[pos + i] += v;
        }
        pos += kLen + (r >> 32) % 100;  // gap2
    }
    if (threadIdx.x == 0 && n_frames) atomicAdd(n_frames, (unsigned long long)k);
}

}  // namespace

int tmk_launch_synth_noise(uint64_t seed, uint64_t total, float sigma, uint64_t n_per_stream, TmkSynthImp imp,
                           float* samples, cudaStream_t st) {
    if (total == 0) return 0;
    const uint64_t threads = (total + 3) / 4;
    noise_kernel<<<uint32_t((threads + 255) / 256), 256, 0, st>>>(seed, total, sigma, n_per_stream, imp, samples);
    return 1;
}

int tmk_launch_ask_synth(const float* tab, uint64_t seed, uint32_t n_streams, uint64_t n_per_stream, float* samples,
                         unsigned long long* n_frames, cudaStream_t st) {
    if (n_streams == 0) return 0;
    ask_synth_kernel<<<n_streams, 256, 0, st>>>(tab, seed, n_per_stream, samples, n_frames);
    return 1;
}

int tmk_launch_synth_plan(const TmkPattern* pattern, int32_t local_addr, uint64_t seed, uint32_t n_streams,
                          uint64_t n_per_stream, TmkSynthImp imp, uint32_t* counts, const uint64_t* base,
                          tmk_frame_rec* frames, uint64_t frame_cap, cudaStream_t st) {
    if (n_streams == 0) return 0;
    plan_kernel<<<(n_streams + 127) / 128, 128, 0, st>>>(pattern, local_addr, seed, n_streams, n_per_stream, imp, counts,
                                                        base, frames, frame_cap);
    return 1;
}

int tmk_launch_synth_frames(const TmkPattern* pattern, uint64_t seed, uint64_t n_per_stream, float amplitude,
                            TmkSynthImp imp, tmk_frame_rec* frames, uint64_t n_frames, float* samples, cudaStream_t st) {
    if (n_frames == 0) return 0;
    frames_kernel<<<uint32_t(n_frames), 128, 0, st>>>(pattern, seed, n_per_stream, amplitude, imp, frames, samples);
    return 1;
}
