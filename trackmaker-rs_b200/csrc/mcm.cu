// mcm.cu — multi-carrier modem: FSK and OFDM demodulation on a batched shared-memory FFT, symbol decision, bit
// packing, CRC-8 (north_star stages 3-4, BASELINE config 3).
//
// THE REFERENCE HAS NO CODE FOR THIS PATH (no FFT / FSK / OFDM in any .rs file; docs/proj/report.md:177-319
// is prose about an amodem-style multi-carrier design). The specification is this build's own — DESIGN.md §3.8,
// stated sequentially in oracle/mcm_modem.hpp — and its results are "parity unpinned by the reference". From the
// reference: the chirp preamble and its synchroniser (examples/ask.rs:586-605, :693-723; ask.cu), MSB-first bit
// packing and CRC-8 (src/phy/crc.rs:7-22,41-60).
//
//   mcm_demod_kernel     one warp per locked frame (the records of the chirp synchroniser, left on the device by
//                        tmk_ask_sync_device): the symbols' samples -> shared memory -> Stockham radix-4 FFT
//                        (256 points, 4 passes, two butterflies per lane; 64 points, 3 passes, one butterfly per lane
//                        of a half-warp, two symbols at a time) -> OFDM: one-tap equalisation by the training symbol
//                        and QPSK sign decisions, a lane per carrier; FSK: strongest of 16 bins -> bytes -> length
//                        rule, CRC-8 -> record + payload.
//   mcm_modulate_kernel  TX, one CTA per frame.   mcm_synth_kernel  capture generator (measurement harness).
// Every f32 operation is rounded separately (__fadd_rn / __fmul_rn: no FMA contraction) in the order of the
// specification, tables come from the host in double -> f32, so spectra, decisions, bytes and verdicts equal the
// oracle's bit for bit.
// Bound: the demodulation touches 4752 of every ~5200+ samples once more (L2 / HBM re-read of the frame bodies) and is
// ~5 K instructions per lane and frame: it is small next to the chirp filter, which reads every sample.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "engine.hpp"

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kPre = TMK_ASK_PREAMBLE, kBody = TMK_MCM_BODY_SAMPLES, kLead = 16;
constexpr int kN = 256, kCp = 64, kCarriers = 32, kBin0 = 16, kDataSyms = 13;
constexpr int kNf = 64, kCpf = 16, kTones = 16, kToneBin0 = 4, kFskSyms = 56;
constexpr int kBytesOfdm = kDataSyms * kCarriers * 2 / 8, kBytesFsk = kFskSyms * 4 / 8;  // 104, 28
constexpr float kGainData = 0.088388346f, kGainTrain = 0.125f, kGainFsk = 0.70710678f;
// table layout (floats): W256 (256 x re,im) | W64 (64 x re,im) | C256 | S256 | S64 | training (32 x u32)
constexpr int kTabW256 = 0, kTabW64 = 512, kTabC = 640, kTabS = 896, kTabS64 = 1152, kTabTrain = 1216, kTabFloats = 1248;

__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float2 cmul(float2 w, float2 v) {
    return make_float2(fsub(fmul(w.x, v.x), fmul(w.y, v.y)), fadd(fmul(w.x, v.y), fmul(w.y, v.x)));
}

// one radix-4 decimation-in-frequency butterfly of the Stockham pass (sequence length n, stride s) of an N-point FFT
template <int N>
__device__ __forceinline__ void butterfly(const float2* __restrict__ src, float2* __restrict__ dst, const float2* __restrict__ W,
                                          int n, int s, int t) {
    const int n1 = n >> 2;
    const int p = t / s, q = t - p * s;
    const float2 a = src[q + s * p], b = src[q + s * (p + n1)], c = src[q + s * (p + 2 * n1)], d = src[q + s * (p + 3 * n1)];
    const float2 apc = make_float2(fadd(a.x, c.x), fadd(a.y, c.y)), amc = make_float2(fsub(a.x, c.x), fsub(a.y, c.y));
    const float2 bpd = make_float2(fadd(b.x, d.x), fadd(b.y, d.y)), bmd = make_float2(fsub(b.x, d.x), fsub(b.y, d.y));
    const float2 y0 = make_float2(fadd(apc.x, bpd.x), fadd(apc.y, bpd.y));
    float2 y1 = make_float2(fadd(amc.x, bmd.y), fsub(amc.y, bmd.x));
    float2 y2 = make_float2(fsub(apc.x, bpd.x), fsub(apc.y, bpd.y));
    float2 y3 = make_float2(fsub(amc.x, bmd.y), fadd(amc.y, bmd.x));
    if (p) {
        const int st = N / n;
        y1 = cmul(W[p * st], y1);
        y2 = cmul(W[2 * p * st], y2);
        y3 = cmul(W[3 * p * st], y3);
    }
    dst[q + s * (4 * p)] = y0;
    dst[q + s * (4 * p + 1)] = y1;
    dst[q + s * (4 * p + 2)] = y2;
    dst[q + s * (4 * p + 3)] = y3;
}

// 256-point FFT by one warp: input in a, result in a (4 passes: a -> b -> a -> b -> a)
__device__ __forceinline__ void fft256_warp(float2* a, float2* b, const float2* __restrict__ W, int lane) {
    float2* src = a;
    float2* dst = b;
#pragma unroll
    for (int n = kN, s = 1; n >= 4; n >>= 2, s <<= 2) {
        butterfly<kN>(src, dst, W, n, s, lane);
        butterfly<kN>(src, dst, W, n, s, lane + 32);
        __syncwarp();
        float2* t = src;
        src = dst;
        dst = t;
    }
}
// 64-point FFT by 16 lanes (l16): input in a, result in b (3 passes: a -> b -> a -> b); both half-warps call it
__device__ __forceinline__ void fft64_half(float2* a, float2* b, const float2* __restrict__ W, int l16) {
    float2* src = a;
    float2* dst = b;
#pragma unroll
    for (int n = kNf, s = 1; n >= 4; n >>= 2, s <<= 2) {
        butterfly<kNf>(src, dst, W, n, s, l16);
        __syncwarp();
        float2* t = src;
        src = dst;
        dst = t;
    }
}

__device__ __forceinline__ uint8_t crc8_bytes(const uint8_t* d, uint32_t n) {  // crc.rs:7-22 (poly 0x07, init 0)
    uint32_t crc = 0;
    for (uint32_t i = 0; i < n; ++i) {
        crc ^= d[i];
        for (int b = 0; b < 8; ++b) crc = (crc & 0x80u) ? ((crc << 1) ^ 0x07u) & 0xffu : (crc << 1) & 0xffu;
    }
    return uint8_t(crc);
}

constexpr int kDemodWarps = 4;

// the bytes of the body that starts at x[0] (x = capture + start_index + 1), into bytes[] (shared, 104 / 28)
__device__ __forceinline__ void demod_body(int mode, const float* __restrict__ x, const float* __restrict__ tab, float2* a, float2* b,
                                           uint8_t* bytes, int lane) {
    if (mode == TMK_MCM_OFDM) {
        const float2* W = reinterpret_cast<const float2*>(tab + kTabW256);
        const uint32_t tr = reinterpret_cast<const uint32_t*>(tab + kTabTrain)[lane];
        float2 R = make_float2(0.f, 0.f);
        for (int sym = 0; sym <= kDataSyms; ++sym) {
            const float* __restrict__ sp = x + kLead + sym * (kN + kCp) + kCp;
#pragma unroll
            for (int i = 0; i < kN / 32; ++i) a[lane + 32 * i] = make_float2(sp[lane + 32 * i], 0.0f);
            __syncwarp();
            fft256_warp(a, b, W, lane);
            const float2 y = a[kBin0 + lane];
            __syncwarp();
            if (sym == 0) {
                R = y;
                continue;
            }
            const float zr = fadd(fmul(y.x, R.x), fmul(y.y, R.y));  // y conj(R)
            const float zi = fsub(fmul(y.y, R.x), fmul(y.x, R.y));
            float er, ei;                                           // times the training point 1, j, -1, -j
            switch (tr) {
                case 0: er = zr; ei = zi; break;
                case 1: er = -zi; ei = zr; break;
                case 2: er = -zr; ei = -zi; break;
                default: er = zi; ei = -zr; break;
            }
            const uint32_t m0 = __ballot_sync(kFull, er < 0.0f), m1 = __ballot_sync(kFull, ei < 0.0f);
            if (lane < 8) {  // byte `lane` of the symbol: carriers 4 lane .. 4 lane + 3, (b0, b1) each, MSB first
                uint32_t v = 0;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int i = 4 * lane + c;
                    v |= ((m0 >> i) & 1u) << (7 - 2 * c);
                    v |= ((m1 >> i) & 1u) << (6 - 2 * c);
                }
                bytes[(sym - 1) * 8 + lane] = uint8_t(v);
            }
        }
    } else {
        const float2* W = reinterpret_cast<const float2*>(tab + kTabW64);
        const int hw = lane >> 4, l16 = lane & 15;
        float2* ah = a + hw * kNf;
        float2* bh = b + hw * kNf;
        for (int it = 0; it < kFskSyms / 2; ++it) {
            const float* __restrict__ sp = x + kLead + (2 * it + hw) * (kNf + kCpf) + kCpf;
#pragma unroll
            for (int i = 0; i < kNf / 16; ++i) ah[l16 + 16 * i] = make_float2(sp[l16 + 16 * i], 0.0f);
            __syncwarp();
            fft64_half(ah, bh, W, l16);
            const float2 v = bh[kToneBin0 + l16];
            __syncwarp();
            const float e = fadd(fmul(v.x, v.x), fmul(v.y, v.y));
            float eb = 0.0f;
            int best = 0;
#pragma unroll
            for (int m = 0; m < kTones; ++m) {  // the first strongest tone, sequentially like the specification
                const float em = __shfl_sync(kFull, e, m, 16);
                if (m == 0 || em > eb) {
                    eb = em;
                    best = m;
                }
            }
            const int other = __shfl_sync(kFull, best, 16);
            if (lane == 0) bytes[it] = uint8_t((best << 4) | other);
        }
    }
    __syncwarp();
}

struct McmCounters {
    unsigned long long n_payload;
};

__global__ void __launch_bounds__(kDemodWarps * 32)
mcm_demod_kernel(int mode, const float* __restrict__ samples, const uint64_t* __restrict__ stream_off,
                 const tmk_ask_rec* __restrict__ sync, uint64_t n, const float* __restrict__ tab, tmk_mcm_rec* __restrict__ out,
                 uint8_t* __restrict__ payload, uint64_t payload_cap, McmCounters* __restrict__ counters) {
    __shared__ float2 buf_a[kDemodWarps][kN], buf_b[kDemodWarps][kN];
    __shared__ uint8_t bytes_s[kDemodWarps][kBytesOfdm + 8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* bytes = bytes_s[warp];
    for (uint64_t r = uint64_t(blockIdx.x) * kDemodWarps + warp; r < n; r += uint64_t(gridDim.x) * kDemodWarps) {
        const tmk_ask_rec sr = sync[r];
        const float* __restrict__ x = samples + stream_off[sr.stream] + sr.start_index + 1;
        demod_body(mode, x, tab, buf_a[warp], buf_b[warp], bytes, lane);
        const uint32_t len = bytes[1];
        const uint32_t cap = mode == TMK_MCM_OFDM ? TMK_MCM_MAX_PAYLOAD_OFDM : TMK_MCM_MAX_PAYLOAD_FSK;
        uint32_t status = TMK_FRAME_BAD_LEN, crc = 0;
        unsigned long long poff = 0;
        if (len <= cap) {
            if (lane == 0) {
                crc = bytes[2 + len];
                status = crc8_bytes(bytes + 2, len) == crc ? TMK_FRAME_OK : TMK_FRAME_CRC_FAIL;
                if (len) poff = atomicAdd(&counters->n_payload, (unsigned long long)len);
            }
            poff = __shfl_sync(kFull, poff, 0);
            if (poff + len <= payload_cap)
                for (uint32_t i = lane; i < len; i += 32) payload[poff + i] = bytes[2 + i];
        }
        if (lane == 0) {
            tmk_mcm_rec o;
            o.stream = sr.stream;
            o.status = status;
            o.start_index = sr.start_index;
            o.payload_off = poff;
            o.sync_power = sr.sync_power;
            o.seq = bytes[0];
            o.len = uint8_t(len);
            o.crc = uint8_t(crc);
            o.mode = uint8_t(mode);
            out[r] = o;
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------- TX
struct FrameSyms {  // the symbols of one frame, in shared memory
    float sr[kDataSyms + 1][kCarriers], si[kDataSyms + 1][kCarriers];
    uint8_t tone[kFskSyms];
    uint8_t bytes[kBytesOfdm + 8];
};

// bytes -> symbols (all threads of the CTA; bytes already in f.bytes, behind a barrier)
__device__ __forceinline__ void build_symbols(int mode, FrameSyms& f, const float* __restrict__ tab) {
    auto bit = [&](int j) { return (f.bytes[j >> 3] >> (7 - (j & 7))) & 1; };
    if (mode == TMK_MCM_OFDM) {
        const uint32_t* tr = reinterpret_cast<const uint32_t*>(tab + kTabTrain);
        for (int k = threadIdx.x; k < (kDataSyms + 1) * kCarriers; k += blockDim.x) {
            const int sym = k / kCarriers, i = k % kCarriers;
            if (sym == 0) {
                const uint32_t t = tr[i];
                const float re = t == 0 ? 1.f : t == 2 ? -1.f : 0.f, im = t == 1 ? 1.f : t == 3 ? -1.f : 0.f;
                f.sr[0][i] = fmul(kGainTrain, re);
                f.si[0][i] = fmul(kGainTrain, im);
            } else {
                const int j = ((sym - 1) * kCarriers + i) * 2;
                f.sr[sym][i] = bit(j) ? -kGainData : kGainData;
                f.si[sym][i] = bit(j + 1) ? -kGainData : kGainData;
            }
        }
    } else {
        for (int sym = threadIdx.x; sym < kFskSyms; sym += blockDim.x)
            f.tone[sym] = uint8_t((bit(4 * sym) << 3) | (bit(4 * sym + 1) << 2) | (bit(4 * sym + 2) << 1) | bit(4 * sym + 3));
    }
}

// body sample b (0 .. 4751) of the frame whose symbols are in f
__device__ __forceinline__ float body_sample(int mode, const FrameSyms& f, const float* __restrict__ tab, int b) {
    if (mode == TMK_MCM_OFDM) {
        const int k = b - kLead;
        if (k < 0 || k >= (kDataSyms + 1) * (kN + kCp)) return 0.0f;
        const int sym = k / (kN + kCp), n = k - sym * (kN + kCp);
        const int t = (n + kN - kCp) & (kN - 1);
        const float* __restrict__ C = tab + kTabC;
        const float* __restrict__ S = tab + kTabS;
        float acc = 0.0f;
#pragma unroll 4
        for (int i = 0; i < kCarriers; ++i) {
            const int m = ((kBin0 + i) * t) & (kN - 1);
            acc = fadd(acc, fmul(f.sr[sym][i], C[m]));
            acc = fsub(acc, fmul(f.si[sym][i], S[m]));
        }
        return acc;
    }
    const int k = b - kLead;
    if (k < 0 || k >= kFskSyms * (kNf + kCpf)) return 0.0f;
    const int sym = k / (kNf + kCpf), n = k - sym * (kNf + kCpf);
    const int t = (n + kNf - kCpf) & (kNf - 1);
    return fmul(kGainFsk, tab[kTabS64 + (((kToneBin0 + f.tone[sym]) * t) & (kNf - 1))]);
}

__device__ __forceinline__ void build_bytes(int mode, FrameSyms& f, uint8_t seq, const uint8_t* __restrict__ payload, uint32_t len) {
    const int nb = mode == TMK_MCM_OFDM ? kBytesOfdm : kBytesFsk;
    for (int i = threadIdx.x; i < nb; i += blockDim.x) f.bytes[i] = 0;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < len; i += blockDim.x) f.bytes[2 + i] = payload[i];
    __syncthreads();
    if (threadIdx.x == 0) {
        f.bytes[0] = seq;
        f.bytes[1] = uint8_t(len);
        f.bytes[2 + len] = crc8_bytes(f.bytes + 2, len);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256)
mcm_modulate_kernel(int mode, const float* __restrict__ tab, const float* __restrict__ chirp, const uint8_t* __restrict__ payload,
                    const uint32_t* __restrict__ pay_off, const uint8_t* __restrict__ seq, const uint64_t* __restrict__ frame_off,
                    const uint32_t* __restrict__ gap1, const uint32_t* __restrict__ gap2, float* __restrict__ out) {
    __shared__ FrameSyms f;
    const uint32_t fr = blockIdx.x;
    build_bytes(mode, f, seq[fr], payload + pay_off[fr], pay_off[fr + 1] - pay_off[fr]);
    build_symbols(mode, f, tab);
    __syncthreads();
    float* o = out + frame_off[fr];
    const uint32_t g1 = gap1[fr], g2 = gap2[fr];
    for (uint32_t i = threadIdx.x; i < g1; i += blockDim.x) o[i] = 0.0f;
    o += g1;
    for (uint32_t i = threadIdx.x; i < kPre; i += blockDim.x) o[i] = chirp[i];
    o += kPre;
    for (int b = threadIdx.x; b < kBody; b += blockDim.x) o[b] = body_sample(mode, f, tab, b);
    o += kBody;
    for (uint32_t i = threadIdx.x; i < g2; i += blockDim.x) o[i] = 0.0f;
}

// ---------------------------------------------------------------- capture generator (measurement harness)
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// one CTA per stream: frames (hashed payloads, gaps 100 + U[0, 400)) laid one after the other, each rendered into
// shared memory and added onto the noise through the channel h = [h0, h1 @ t1, h2 @ t2]
__global__ void __launch_bounds__(256)
mcm_synth_kernel(int mode, const float* __restrict__ tab, const float* __restrict__ chirp, uint64_t seed, uint64_t n_per_stream,
                 int multipath, float* __restrict__ samples, unsigned long long* __restrict__ n_frames) {
    __shared__ FrameSyms f;
    __shared__ float clean[kPre + kBody];
    __shared__ uint8_t pay[TMK_MCM_MAX_PAYLOAD_OFDM];
    const uint32_t s = blockIdx.x;
    float* __restrict__ o = samples + uint64_t(s) * n_per_stream;
    constexpr uint32_t kLen = kPre + kBody;
    const uint64_t hs = mix64(seed ^ (uint64_t(s) << 24) ^ 0x3C3ull);
    const int t1 = 2 + int((hs >> 48) % 11), t2 = 13 + int((hs >> 56) % 28);
    const float h0 = multipath ? 0.92747675f : 1.0f, h1 = multipath ? 0.32461686f : 0.0f, h2 = multipath ? -0.18549535f : 0.0f;
    const uint32_t cap = mode == TMK_MCM_OFDM ? TMK_MCM_MAX_PAYLOAD_OFDM : TMK_MCM_MAX_PAYLOAD_FSK;
    uint64_t st = mix64(seed ^ (uint64_t(s) << 20) ^ 0x3C3Dull);
    uint64_t pos = 0;
    uint32_t k = 0;
    for (;; ++k) {
        st = mix64(st);
        const uint64_t r = st;
        pos += 100 + r % 400;
        if (pos + kLen + 64 > n_per_stream) break;
        const uint32_t len = 1 + uint32_t((r >> 16) % cap);
        for (uint32_t i = threadIdx.x; i < len; i += blockDim.x) pay[i] = uint8_t(mix64(r ^ (uint64_t(i) << 32)) >> 11);
        __syncthreads();
        build_bytes(mode, f, uint8_t(k), pay, len);
        build_symbols(mode, f, tab);
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < kLen; i += blockDim.x)
            clean[i] = i < uint32_t(kPre) ? chirp[i] : body_sample(mode, f, tab, int(i) - kPre);
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < kLen + uint32_t(multipath ? t2 : 0); i += blockDim.x) {
            float v = i < kLen ? h0 * clean[i] : 0.0f;
            if (multipath) {
                if (i >= uint32_t(t1) && i - t1 < kLen) v += h1 * clean[i - t1];
                if (i >= uint32_t(t2) && i - t2 < kLen) v += h2 * clean[i - t2];
            }
            o[pos + i] += v;
        }
        __syncthreads();
        pos += kLen;
    }
    if (threadIdx.x == 0 && n_frames) atomicAdd(n_frames, (unsigned long long)k);
}

// parity probe of the FFT: one warp per row
__global__ void __launch_bounds__(kDemodWarps * 32)
mcm_fft_probe_kernel(int n, const float* __restrict__ rows, uint32_t n_ffts, const float* __restrict__ tab, float2* __restrict__ out) {
    __shared__ float2 buf_a[kDemodWarps][kN], buf_b[kDemodWarps][kN];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float2* a = buf_a[warp];
    float2* b = buf_b[warp];
    for (uint32_t r = blockIdx.x * kDemodWarps + warp; r < n_ffts; r += gridDim.x * kDemodWarps) {
        const float* __restrict__ x = rows + uint64_t(r) * n;
        const float2* res;
        if (n == kN) {
            for (int i = lane; i < kN; i += 32) a[i] = make_float2(x[i], 0.0f);
            __syncwarp();
            fft256_warp(a, b, reinterpret_cast<const float2*>(tab + kTabW256), lane);
            res = a;
        } else {  // both half-warps run the same 64-point transform on their own copy (the demodulator's code path)
            float2* ah = a + (lane >> 4) * kNf;
            float2* bh = b + (lane >> 4) * kNf;
            for (int i = lane & 15; i < kNf; i += 16) ah[i] = make_float2(x[i], 0.0f);
            __syncwarp();
            fft64_half(ah, bh, reinterpret_cast<const float2*>(tab + kTabW64), lane & 15);
            res = b;
        }
        for (int i = lane; i < n; i += 32) out[uint64_t(r) * n + i] = res[i];
        __syncwarp();
    }
}

int ensure_mcm_tables(tmk_engine* e) {
    if (e->mcm_ready) return TMK_OK;
    std::vector<float> tab(kTabFloats, 0.0f);
    const double two_pi = 6.283185307179586476925286766559;
    for (int t = 0; t < kN; ++t) {
        const double a = two_pi * double(t) / double(kN);
        tab[kTabW256 + 2 * t] = float(std::cos(a));
        tab[kTabW256 + 2 * t + 1] = float(-std::sin(a));
        tab[kTabC + t] = float(std::cos(a));
        tab[kTabS + t] = float(std::sin(a));
    }
    for (int t = 0; t < kNf; ++t) {
        const double a = two_pi * double(t) / double(kNf);
        tab[kTabW64 + 2 * t] = float(std::cos(a));
        tab[kTabW64 + 2 * t + 1] = float(-std::sin(a));
        tab[kTabS64 + t] = float(std::sin(a));
    }
    uint32_t st = 0x1234u;
    for (int i = 0; i < kCarriers; ++i) {
        st = st * 1103515245u + 12345u;
        const uint32_t v = (st >> 16) & 3u;
        std::memcpy(&tab[kTabTrain + i], &v, sizeof v);
    }
    CK(e->d_mcm_tab.reserve(tab.size() * sizeof(float)));
    CK(cudaMemcpyAsync(e->d_mcm_tab.p, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    e->mcm_ready = true;
    return TMK_OK;
}

bool bad_mode(int32_t mode) { return mode != TMK_MCM_FSK && mode != TMK_MCM_OFDM; }

}  // namespace

extern "C" {

uint64_t tmk_mcm_modulated_len(uint32_t n_frames, const uint32_t* gap1, const uint32_t* gap2) {
    uint64_t total = 0;
    for (uint32_t f = 0; f < n_frames; ++f) total += uint64_t(gap1 ? gap1[f] : 0) + kPre + kBody + (gap2 ? gap2[f] : 0);
    return total;
}

int tmk_mcm_modulate(tmk_engine* e, int32_t mode, const uint8_t* payload, const uint32_t* payload_off, const uint8_t* seq,
                     uint32_t n_frames, const uint32_t* gap1, const uint32_t* gap2, float* out_samples, uint64_t out_cap,
                     uint64_t* out_len) {
    if (!e || !out_len || ((!payload_off || !seq || !gap1 || !gap2) && n_frames))
        return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_modulate: null argument");
    if (bad_mode(mode)) return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_modulate: mode must be TMK_MCM_FSK or TMK_MCM_OFDM");
    CK(cudaSetDevice(e->device));
    const uint32_t cap = mode == TMK_MCM_OFDM ? TMK_MCM_MAX_PAYLOAD_OFDM : TMK_MCM_MAX_PAYLOAD_FSK;
    std::vector<uint64_t> off(n_frames + 1, 0);
    for (uint32_t f = 0; f < n_frames; ++f) {
        if (payload_off[f + 1] < payload_off[f] || payload_off[f + 1] - payload_off[f] > cap)
            return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_modulate: frame " + std::to_string(f) + " exceeds the payload capacity of the mode");
        off[f + 1] = off[f] + gap1[f] + kPre + kBody + gap2[f];
    }
    const uint64_t total = off[n_frames];
    *out_len = total;
    if (total > out_cap) return tmk_fail(e, TMK_E_CAPACITY, "tmk_mcm_modulate: output buffer too small");
    if (total == 0) return TMK_OK;
    if (!out_samples || (!payload && payload_off[n_frames])) return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_modulate: null buffer");
    int rc = tmk_ask_ensure_tables(e);
    if (rc != TMK_OK) return rc;
    rc = ensure_mcm_tables(e);
    if (rc != TMK_OK) return rc;
    const size_t ob = sizeof(uint64_t) * (n_frames + 1), gb = sizeof(uint32_t) * n_frames, pob = sizeof(uint32_t) * (n_frames + 1);
    const size_t sb = (size_t(n_frames) + 7) / 8 * 8, pb = size_t(payload_off[n_frames]);
    CK(e->d_misc.reserve(ob + 2 * gb + pob + sb + pb + 64));
    char* dm = static_cast<char*>(e->d_misc.p);
    CK(cudaMemcpyAsync(dm, off.data(), ob, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(dm + ob, gap1, gb, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(dm + ob + gb, gap2, gb, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(dm + ob + 2 * gb, payload_off, pob, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(dm + ob + 2 * gb + pob, seq, n_frames, cudaMemcpyHostToDevice, e->stream));
    if (pb) CK(cudaMemcpyAsync(dm + ob + 2 * gb + pob + sb, payload, pb, cudaMemcpyHostToDevice, e->stream));
    float* d_out = out_samples;
    const bool dev_out = tmk_is_device_ptr(out_samples);
    if (!dev_out) {
        CK(e->d_samples.reserve(total * sizeof(float)));
        d_out = static_cast<float*>(e->d_samples.p);
    }
    mcm_modulate_kernel<<<n_frames, 256, 0, e->stream>>>(
        mode, static_cast<const float*>(e->d_mcm_tab.p), static_cast<const float*>(e->d_ask_tab.p),
        reinterpret_cast<const uint8_t*>(dm + ob + 2 * gb + pob + sb), reinterpret_cast<const uint32_t*>(dm + ob + 2 * gb),
        reinterpret_cast<const uint8_t*>(dm + ob + 2 * gb + pob), reinterpret_cast<const uint64_t*>(dm),
        reinterpret_cast<const uint32_t*>(dm + ob), reinterpret_cast<const uint32_t*>(dm + ob + gb), d_out);
    e->launches += 1;
    CK(cudaGetLastError());
    if (!dev_out) CK(cudaMemcpyAsync(out_samples, d_out, total * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_mcm_decode_batch(tmk_engine* e, int32_t mode, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                         uint32_t n_streams, tmk_mcm_rec* recs, uint64_t rec_cap, uint64_t* n_recs, uint8_t* payload,
                         uint64_t payload_cap, uint64_t* payload_len) {
    if (!e || (!samples && total_samples) || (!streams && n_streams) || !n_recs || !payload_len || (!recs && rec_cap) ||
        (!payload && payload_cap))
        return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_decode_batch: null argument");
    if (bad_mode(mode)) return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_decode_batch: mode must be TMK_MCM_FSK or TMK_MCM_OFDM");
    *n_recs = 0;
    *payload_len = 0;
    if (n_streams == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    int rc = ensure_mcm_tables(e);
    if (rc != TMK_OK) return rc;
    // the chirp synchroniser of examples/ask.rs finds the frames (records stay on the device)
    const float* dx = nullptr;
    uint64_t n = 0, dev_cap = 0;
    rc = tmk_ask_sync_device(e, samples, total_samples, streams, n_streams, &dx, &n, &dev_cap, false);  // no BPSK demodulation
    *n_recs = n;
    if (rc != TMK_OK) return rc;
    if (n > rec_cap) return tmk_fail(e, TMK_E_CAPACITY, "tmk_mcm_decode_batch: record buffer too small");
    e->mcm_demod_ms = 0.0f;
    if (n == 0) return TMK_OK;
    std::vector<uint64_t> off(n_streams);
    for (uint32_t s = 0; s < n_streams; ++s) off[s] = streams[s].offset;
    const uint64_t dev_pay_cap = std::min<uint64_t>(payload_cap, n * TMK_MCM_MAX_PAYLOAD_OFDM);
    CK(e->d_mcm_off.reserve(sizeof(uint64_t) * n_streams + sizeof(McmCounters) + 64));
    CK(e->d_mcm_recs.reserve(sizeof(tmk_mcm_rec) * n));
    CK(e->d_payload.reserve(std::max<uint64_t>(1, dev_pay_cap)));
    McmCounters* dc = reinterpret_cast<McmCounters*>(static_cast<char*>(e->d_mcm_off.p) + (sizeof(uint64_t) * n_streams + 15) / 16 * 16);
    CK(cudaMemcpyAsync(e->d_mcm_off.p, off.data(), sizeof(uint64_t) * n_streams, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemsetAsync(dc, 0, sizeof(McmCounters), e->stream));
    CK(cudaEventRecord(e->ev[3], e->stream));
    static int max_ctas = 0;
    if (max_ctas == 0) {
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mcm_demod_kernel, kDemodWarps * 32, 0);
        max_ctas = sms * (per_sm > 0 ? per_sm : 1);
    }
    const uint32_t grid = uint32_t(std::min<uint64_t>((n + kDemodWarps - 1) / kDemodWarps, uint64_t(max_ctas)));
    mcm_demod_kernel<<<grid, kDemodWarps * 32, 0, e->stream>>>(
        mode, dx, static_cast<const uint64_t*>(e->d_mcm_off.p), static_cast<const tmk_ask_rec*>(e->d_ask_recs.p), n,
        static_cast<const float*>(e->d_mcm_tab.p), static_cast<tmk_mcm_rec*>(e->d_mcm_recs.p),
        static_cast<uint8_t*>(e->d_payload.p), dev_pay_cap, dc);
    e->launches += 1;
    CK(cudaGetLastError());
    CK(cudaEventRecord(e->ev[4], e->stream));
    McmCounters hc;
    CK(cudaMemcpyAsync(&hc, dc, sizeof hc, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(recs, e->d_mcm_recs.p, n * sizeof(tmk_mcm_rec), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    cudaEventElapsedTime(&e->mcm_demod_ms, e->ev[3], e->ev[4]);
    *payload_len = hc.n_payload;
    if (hc.n_payload > payload_cap) return tmk_fail(e, TMK_E_CAPACITY, "tmk_mcm_decode_batch: payload buffer too small");
    if (hc.n_payload) {
        CK(cudaMemcpyAsync(payload, e->d_payload.p, hc.n_payload, cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
    }
    // (the synchroniser's records are in (capture, time) order already, and record i here is its record i)
    return TMK_OK;
}

int tmk_mcm_synth_batch(tmk_engine* e, int32_t mode, uint64_t seed, uint32_t n_streams, uint64_t n_per_stream, float snr_db,
                        int32_t multipath, float* samples, uint64_t* n_frames) {
    if (!e || !samples || !n_frames) return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_synth_batch: null argument");
    if (bad_mode(mode)) return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_synth_batch: mode must be TMK_MCM_FSK or TMK_MCM_OFDM");
    if (n_per_stream % 4 != 0 || n_per_stream >= (1ull << 32))
        return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_synth_batch: n_per_stream must be a multiple of 4 below 2^32");
    *n_frames = 0;
    if (n_streams == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    int rc = tmk_ask_ensure_tables(e);
    if (rc != TMK_OK) return rc;
    rc = ensure_mcm_tables(e);
    if (rc != TMK_OK) return rc;
    const uint64_t total = uint64_t(n_streams) * n_per_stream;
    float* d_x = samples;
    const bool dev_out = tmk_is_device_ptr(samples);
    if (!dev_out) {
        CK(e->d_samples.reserve(total * sizeof(float)));
        d_x = static_cast<float*>(e->d_samples.p);
    }
    TmkSynthImp imp;
    std::memset(&imp, 0, sizeof imp);
    const float sigma = snr_db >= 100.0f ? 0.0f : 0.5f * std::pow(10.0f, -snr_db / 20.0f);  // body power 0.25
    e->launches += tmk_launch_synth_noise(seed ^ 0x3C3ull, total, sigma, n_per_stream, imp, d_x, e->stream);
    CK(e->d_ask_l2.reserve(4 * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(e->d_ask_l2.p, 0, 4 * sizeof(unsigned long long), e->stream));
    mcm_synth_kernel<<<n_streams, 256, 0, e->stream>>>(mode, static_cast<const float*>(e->d_mcm_tab.p),
                                                       static_cast<const float*>(e->d_ask_tab.p), seed, n_per_stream, multipath ? 1 : 0,
                                                       d_x, static_cast<unsigned long long*>(e->d_ask_l2.p));
    e->launches += 1;
    CK(cudaGetLastError());
    unsigned long long nf = 0;
    CK(cudaMemcpyAsync(&nf, e->d_ask_l2.p, sizeof nf, cudaMemcpyDeviceToHost, e->stream));
    if (!dev_out) CK(cudaMemcpyAsync(samples, d_x, total * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    *n_frames = nf;
    return TMK_OK;
}

int tmk_mcm_fft_probe(tmk_engine* e, uint32_t n, const float* rows, uint32_t n_ffts, float* spectra) {
    if (!e || (!rows && n_ffts) || (!spectra && n_ffts)) return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_fft_probe: null argument");
    if (n != uint32_t(kN) && n != uint32_t(kNf)) return tmk_fail(e, TMK_E_INVALID, "tmk_mcm_fft_probe: n must be 64 or 256");
    if (n_ffts == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    int rc = ensure_mcm_tables(e);
    if (rc != TMK_OK) return rc;
    const size_t in_b = size_t(n_ffts) * n * sizeof(float), out_b = 2 * in_b;
    CK(e->d_misc.reserve(in_b));
    CK(e->d_misc2.reserve(out_b));
    CK(cudaMemcpyAsync(e->d_misc.p, rows, in_b, cudaMemcpyDefault, e->stream));
    mcm_fft_probe_kernel<<<std::min<uint32_t>((n_ffts + kDemodWarps - 1) / kDemodWarps, 148 * 8), kDemodWarps * 32, 0, e->stream>>>(
        int(n), static_cast<const float*>(e->d_misc.p), n_ffts, static_cast<const float*>(e->d_mcm_tab.p),
        static_cast<float2*>(e->d_misc2.p));
    e->launches += 1;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(spectra, e->d_misc2.p, out_b, cudaMemcpyDefault, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_mcm_last_stage_ms(tmk_engine* e, float ms[3]) {
    if (!e || !ms) return TMK_E_INVALID;
    ms[0] = e->ask_ms[0];
    ms[1] = e->ask_ms[1];
    ms[2] = e->mcm_demod_ms;
    return TMK_OK;
}

}  // extern "C"
