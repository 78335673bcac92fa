// ask.cu — the chirp + BPSK modem of the reference's project 1, examples/ask.rs::test_sender_receiver
// (:526-858; live TX :104-262, live RX :264-524), batched over many captures.
//
//   ask_scan_fft_kernel  FILTER over every sample: the 440-tap chirp correlation / 200 as a fast correlation
//                     (2048-point Stockham FFT in shared memory, two tiles packed as re / im) and the EMA power
//                     (affine block scan), marks every 32-sample word in which
//                     `sync_power > 2 power && sync_power > 0.05` (ask.rs:711-714) could hold, with margins that
//                     cover the FFT's error bound and the re-association error.
//   ask_scan_kernel   the direct-form version of the same filter (fp32 FMA, register-blocked out of shared
//                     memory), kept behind TMK_ASK_DIRECT=1 for A/B.
//   ask_walk_kernel   one warp per capture replays the reference's state machine over the marked words
//                     only, in the reference's f32 operation order: sequential 440-term dot product
//                     (ask.rs:704-709), the exact EMA chain (:696-697, carried over every sample), running
//                     maximum with the 200-sample hold-off (:711-723), the dropped sample start_index+201,
//                     the zeroed FIFO after a frame, carrier mixing + 11-tap moving average + integrate and
//                     dump (:733-761), id check (:764-771).
//   ask_modulate_kernel  TX (:608-637).
// Tables (preamble, carrier) are computed once on the host with sinf in the reference's operation order
// (ask.rs:574-605) and uploaded, so device sin never enters.
//
// Bound: the FFT filter is shared-memory bound (~120 flop per sample, 12 passes over 16 KB per 2048 positions);
// the direct form was FP32-pipe bound at 880 flop per 4-byte sample.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "engine.hpp"

namespace {

constexpr int kPre = TMK_ASK_PREAMBLE;
constexpr int kFrame = TMK_ASK_FRAME_SAMPLES;
constexpr int kBits = TMK_ASK_FRAME_BITS;
constexpr unsigned kFull = 0xffffffffu;
constexpr float kC = 0.984375f;     // 1.0 - 1.0 / 64.0   (exact)
constexpr float kInv64 = 0.015625f;  // x * x / 64.0: scaling by a power of two, same rounding as the divide

struct AskStream {
    uint64_t offset, n;
    uint64_t word_base;  // first flag bit of this stream (one bit per 32 samples)
    uint32_t tile_base;  // first scan tile
    uint32_t seg_base;   // first segment of the exact power chain (kSeg samples each)
};

// ---------------------------------------------------------------- TX
__global__ void ask_modulate_kernel(const float* __restrict__ tab, const uint8_t* __restrict__ bits,
                                    const uint64_t* __restrict__ frame_off, const uint32_t* __restrict__ gap1,
                                    const uint32_t* __restrict__ gap2, float* __restrict__ out) {
    const uint32_t f = blockIdx.x;
    float* o = out + frame_off[f];
    const uint32_t g1 = gap1[f], g2 = gap2[f];
    const float* pre = tab;
    const float* car = tab + kPre;
    for (uint32_t i = threadIdx.x; i < g1; i += blockDim.x) o[i] = 0.0f;
    o += g1;
    for (uint32_t i = threadIdx.x; i < kPre; i += blockDim.x) o[i] = pre[i];
    o += kPre;
    for (uint32_t k = threadIdx.x; k < kFrame; k += blockDim.x) {
        const uint32_t j = k / TMK_ASK_SAMPLES_PER_BIT;
        const bool one = j < TMK_ASK_DATA_BITS && bits[f * TMK_ASK_DATA_BITS + j] == 1;
        o[k] = one ? car[k] : -car[k];  // carrier[k] * (+-1.0) is exact
    }
    o += kFrame;
    for (uint32_t i = threadIdx.x; i < g2; i += blockDim.x) o[i] = 0.0f;
}

// ---------------------------------------------------------------- filter
constexpr int kT = 1024;      // positions per tile
constexpr int kThreads = 128; // 8 positions per thread
constexpr int kW = 2048;      // EMA warm-up: 0.984375^2048 ~ 1e-14
constexpr int kXs = kW + kT;  // staged samples: x[t0 - kW .. t0 + kT)

__global__ void __launch_bounds__(kThreads)
ask_scan_kernel(const float* __restrict__ samples, const AskStream* __restrict__ streams, uint32_t n_streams,
                const float* __restrict__ tab, uint32_t* __restrict__ wflag) {
    __shared__ __align__(16) float xs[kXs];
    __shared__ __align__(16) float ps[kT];
    __shared__ __align__(16) float pre[kPre];
    __shared__ float red[kThreads / 32];
    __shared__ float carry[kThreads / 32];
    const uint32_t tile = blockIdx.x;
    uint32_t lo = 0, hi = n_streams - 1;
    while (lo < hi) {
        const uint32_t mid = (lo + hi + 1) >> 1;
        if (streams[mid].tile_base <= tile) lo = mid; else hi = mid - 1;
    }
    const AskStream sd = streams[lo];
    const float* __restrict__ x = samples + sd.offset;
    const int64_t n = int64_t(sd.n);
    const int64_t t0 = int64_t(tile - sd.tile_base) * kT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kPre; i += kThreads) pre[i] = tab[i];
    float xmax = 0.0f;
    for (int i = tid; i < kXs; i += kThreads) {
        const int64_t g = t0 - kW + i;
        const float v = (g >= 0 && g < n) ? x[g] : 0.0f;
        xs[i] = v;
        xmax = fmaxf(xmax, fabsf(v));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) xmax = fmaxf(xmax, __shfl_xor_sync(kFull, xmax, o));
    if (lane == 0) red[warp] = xmax;
    __syncthreads();
    xmax = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
    if (!(xmax <= 3.0e38f)) xmax = 3.0e38f;  // NaN / inf samples: everything in the tile becomes a candidate

    // ---- EMA power p_i = c p_{i-1} + x_i^2 / 64 over the staged range, zero state at its start
    constexpr int kPer = kXs / kThreads;  // 24 consecutive samples per thread
    float q = 0.0f;
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
        const float v = xs[tid * kPer + j];
        q = fmaf(q, kC, v * v * kInv64);
    }
    // inclusive scan of the affine maps p -> c^24 p + q over the threads (c^24 constant per step)
    float v = q, mul = powf(kC, float(kPer));
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const float t = __shfl_up_sync(kFull, v, o);
        if (lane >= o) v = fmaf(t, mul, v);
        mul *= mul;
    }
    if (lane == 31) carry[warp] = v;
    __syncthreads();
    const float c768 = powf(kC, float(kPer * 32));  // one warp's span
    float pin = 0.0f;  // state entering this warp
    for (int w = 0; w < warp; ++w) pin = fmaf(pin, c768, carry[w]);
    // state entering this thread = (inclusive of previous lane) advanced
    float prev = __shfl_up_sync(kFull, v, 1);
    if (lane == 0) prev = 0.0f;
    float p = fmaf(pin, powf(kC, float(kPer * lane)), prev);
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
        const int i = tid * kPer + j;
        const float xv = xs[i];
        p = fmaf(p, kC, xv * xv * kInv64);
        if (i >= kW) ps[i - kW] = p;
    }
    __syncthreads();

    // ---- 440-tap correlation, 8 consecutive positions per thread: position t0 + 8 tid + j uses
    //      x[i - 439 + k] = xs[kW - 439 + 8 tid + j + k]
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.0f;
    const float* xp = xs + (kW - 440) + 8 * tid;  // 16-byte aligned; element [1 + j + k] is the sample of tap k
#pragma unroll 2
    for (int k = 0; k < kPre; k += 4) {
        const float4 pr = *reinterpret_cast<const float4*>(pre + k);
        const float4 a = *reinterpret_cast<const float4*>(xp + k);
        const float4 b = *reinterpret_cast<const float4*>(xp + k + 4);
        const float4 c = *reinterpret_cast<const float4*>(xp + k + 8);
        const float xv[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
        const float pk[4] = {pr.x, pr.y, pr.z, pr.w};
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = fmaf(xv[1 + j + kk], pk[kk], acc[j]);
    }
    // ---- candidate test with margins for the re-association error of both sides
    const float tol = 1.3e-4f * xmax + 1e-7f;  // both sums are within 440 u sum|x pre| / 200 <= 5.8e-5 xmax of the exact value
    uint32_t any = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int64_t i = t0 + 8 * tid + j;
        const float sp = acc[j] * (1.0f / 200.0f) + tol;
        const float pw = ps[8 * tid + j];
        const bool cand = i < n && (!(sp <= 0.05f) && !(sp <= 2.0f * pw * (1.0f - 1e-4f)));  // NaN -> candidate
        any |= cand ? 1u : 0u;
    }
    // 4 threads cover one 32-sample word
    any |= __shfl_xor_sync(kFull, any, 1);
    any |= __shfl_xor_sync(kFull, any, 2);
    if ((lane & 3) == 0 && any) {
        const uint64_t wbit = sd.word_base + uint64_t(t0 + 8 * tid) / 32;
        atomicOr(wflag + (wbit >> 5), 1u << (wbit & 31));
    }
}

// ---------------------------------------------------------------- filter, FFT form (overlap-save in shared memory)
// The 440-tap chirp correlation of the filter as a fast correlation: TWO tiles of kT2 = 1536 positions are packed as the
// real and imaginary part of one 2048-point complex sequence z (the chirp is real, so corr(z, h) = corr(re z, h) +
// i corr(im z, h)), z is transformed by a Stockham autosort FFT in shared memory (five radix-4 passes and one radix-2
// pass, twiddles from a table computed in double on the host), multiplied by conj(H) / (N 200) (the chirp's spectrum,
// also from the host), and transformed back. A tile needs z[0 .. kT2 + 439), the rest of the 2048 points is zero
// padding, so nothing wraps around. ~120 flop per position instead of 880.
// The result differs from the exactly rounded sum by at most (Higham, Accuracy and Stability of Numerical Algorithms,
// §24.1: a radix-2 level with correctly rounded twiddles adds a 2-norm relative error eta <= u + gamma_4 (sqrt 2 + u)
// ~ 6.7 u; a radix-4 pass is two such levels) (5 x 2 + 1) x 6.7 u = 74 u per transform, twice, + 4 u for the rounded
// spectrum and the product: ~ 152 u ~ 9.1e-6, of |H|max / 200 x ||z||_2 (Parseval bounds every element by the 2-norm).
// The tolerance takes 2.0e-5 (2.2 x that), with ||z||_2 measured per CTA, next to the budget of the walk's own
// sequential sum (6.5e-5 xmax, see ask_scan_kernel).
constexpr int kN = 2048;
constexpr int kFT = 256;           // threads: half 0 (tid < 128) owns the tile in re z, half 1 the tile in im z
constexpr float kFftRel = 2.0e-5f;
// EMA warm-up of the FFT filter: the truncated EMA only DROPS non-negative terms (0.984375^1024 ~ 1e-7 of the state
// 1024 samples back), so the power it tests against is never above the exact one: truncation can add candidates,
// never lose one.
constexpr int kW2 = 512;            // (0.984375^512 ~ 3e-4; the 0.05 floor of the test keeps silence out)
constexpr int kT2 = 1536;           // positions per tile of the FFT filter: 1536 + 439 of the 2048 points are data
constexpr int kXs2 = kW2 + kT2;     // 2048 staged samples per tile
constexpr int kZ2 = kT2 + kPre - 1;

struct FftSmem {
    float2 buf[2][kN];
    // twiddles, contiguous per pass (a strided walk through one table is a 16-way bank conflict): radix-4 pass with
    // p = 4^s has w^k, w^2k, w^3k (w = exp(-2 pi i / 4p)) at [4^s - 1 + {0, p, 2p} + k]; the radix-2 pass
    // exp(-2 pi i k / N) at [1023 + k]; one pad entry
    float2 tw[kN];
    float2 hc[kN];  // conj(H) / (N 200)
    __align__(16) float ps[2][kT2];  // (the staged samples xs[2][kXs2] live in buf[1] until the first FFT pass overwrites them)
    float red[kFT / 32], red2[kFT / 32];
    float carry[2][4];
};

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }

// 2048-point Stockham FFT: input in a, scratch b, result in a (six passes). INV: conjugated twiddles, no scaling.
// The forward transform multiplies by the spectrum table `hm` on the way out of its last pass.
template <bool INV>
__device__ __forceinline__ void fft2048(float2* a, float2* b, const float2* __restrict__ tw, const float2* __restrict__ hm, int tid) {
    float2 *in = a, *out = b;
#pragma unroll 1
    for (int p = 1; p <= 256; p <<= 2) {
        const float2* __restrict__ tp = tw + (p - 1);
#pragma unroll
        for (int i = tid; i < kN / 4; i += kFT) {
            const int k = i & (p - 1), j = ((i - k) << 2) + k;
            float2 w1 = tp[k], w2 = tp[p + k], w3 = tp[2 * p + k];
            if (INV) {
                w1.y = -w1.y;
                w2.y = -w2.y;
                w3.y = -w3.y;
            }
            const float2 u0 = in[i], u1 = cmul(in[i + kN / 4], w1), u2 = cmul(in[i + kN / 2], w2), u3 = cmul(in[i + 3 * kN / 4], w3);
            const float2 v0 = cadd(u0, u2), v1 = csub(u0, u2), v2 = cadd(u1, u3), d = csub(u1, u3);
            const float2 v3 = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);  // * (+i) backward, * (-i) forward
            const float2 y0 = cadd(v0, v2), y1 = cadd(v1, v3), y2 = csub(v0, v2), y3 = csub(v1, v3);
            if (p == 1) {  // the four results are neighbours: two 16-byte stores instead of four strided 8-byte ones
                *reinterpret_cast<float4*>(out + j) = make_float4(y0.x, y0.y, y1.x, y1.y);
                *reinterpret_cast<float4*>(out + j + 2) = make_float4(y2.x, y2.y, y3.x, y3.y);
            } else {
                out[j] = y0;
                out[j + p] = y1;
                out[j + 2 * p] = y2;
                out[j + 3 * p] = y3;
            }
        }
        __syncthreads();
        float2* t = in;
        in = out;
        out = t;
    }
    // radix 2, p = 1024 (in == b here, out == a)
#pragma unroll
    for (int i = tid; i < kN / 2; i += kFT) {
        float2 w = tw[1023 + i];
        if (INV) w.y = -w.y;
        const float2 u0 = in[i], u1 = cmul(in[i + kN / 2], w);
        if (!INV) {
            out[i] = cmul(cadd(u0, u1), hm[i]);
            out[i + kN / 2] = cmul(csub(u0, u1), hm[i + kN / 2]);
        } else {
            out[i] = cadd(u0, u1);
            out[i + kN / 2] = csub(u0, u1);
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kFT)
ask_scan_fft_kernel(const float* __restrict__ samples, const AskStream* __restrict__ streams, uint32_t n_streams, uint32_t n_tiles,
                    const float2* __restrict__ twg, const float2* __restrict__ hc, float h_scale, uint32_t* __restrict__ wflag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FftSmem& sm = *reinterpret_cast<FftSmem*>(smem_raw);
    const int tid = threadIdx.x, half = tid >> 7, htid = tid & 127, lane = tid & 31, warp = tid >> 5, hwarp = htid >> 5;
    // persistent CTAs: the tables are staged once, the constants of the EMA scan computed once
    for (int i = tid; i < kN; i += kFT) {
        sm.tw[i] = twg[i];
        sm.hc[i] = hc[i];
    }
    constexpr int kPer = kXs2 / 128;
    const float mul0 = powf(kC, float(kPer)), c768 = powf(kC, float(kPer * 32)), clane = powf(kC, float(kPer * lane));
    const uint32_t n_pairs = (n_tiles + 1) / 2;
    // software pipeline over the CTA's pairs: the samples of the NEXT pair are located and loaded into registers while
    // the transforms of the current one run
    constexpr int kLd = kXs2 / 4 / 128;  // 16-byte loads per thread and tile (stream offsets and t0 - kW2 are multiples of 4)
    struct TileAt {
        AskStream sd;
        int64_t n, t0;
        bool valid;
    };
    auto locate = [&](uint32_t pr) {
        TileAt a;
        const uint32_t tl = 2 * pr + half;
        a.valid = pr < n_pairs && tl < n_tiles;
        a.sd = streams[0];
        a.n = 0;
        a.t0 = 0;
        if (a.valid) {
            uint32_t lo = 0, hi = n_streams - 1;
            while (lo < hi) {
                const uint32_t mid = (lo + hi + 1) >> 1;
                if (streams[mid].tile_base <= tl) lo = mid; else hi = mid - 1;
            }
            a.sd = streams[lo];
            a.n = int64_t(a.sd.n);
            a.t0 = int64_t(tl - a.sd.tile_base) * kT2;
        }
        return a;
    };
    auto fetch = [&](const TileAt& a, float4 (&q)[kLd]) {
        const float* __restrict__ x = samples + a.sd.offset;
#pragma unroll
        for (int r = 0; r < kLd; ++r) {
            const int64_t g = a.t0 - kW2 + 4 * (htid + 128 * r);
            float4 q4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (a.valid && g >= 0 && g + 3 < a.n) {
                q4 = __ldg(reinterpret_cast<const float4*>(x + g));
            } else if (a.valid) {
                if (g >= 0 && g < a.n) q4.x = x[g];
                if (g + 1 >= 0 && g + 1 < a.n) q4.y = x[g + 1];
                if (g + 2 >= 0 && g + 2 < a.n) q4.z = x[g + 2];
                if (g + 3 >= 0 && g + 3 < a.n) q4.w = x[g + 3];
            }
            q[r] = q4;
        }
    };
    TileAt nxt = locate(blockIdx.x);
    float4 qn[kLd];
    fetch(nxt, qn);
#pragma unroll 1
    for (uint32_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
    const TileAt cur = nxt;
    const bool valid = cur.valid;
    const AskStream sd = cur.sd;
    const int64_t n = cur.n, t0 = cur.t0;
    float xmax = 0.0f, ss = 0.0f;
    static_assert(sizeof(float) * 2 * kXs2 <= sizeof(float2) * kN, "the staged samples alias buf[1]");
    float* xs0 = reinterpret_cast<float*>(sm.buf[1]);
    float* xs = xs0 + half * kXs2;
#pragma unroll
    for (int r = 0; r < kLd; ++r) {
        const int i = 4 * (htid + 128 * r);
        const float4 q4 = qn[r];
        *reinterpret_cast<float4*>(xs + i) = q4;
        xmax = fmaxf(fmaxf(xmax, fmaxf(fabsf(q4.x), fabsf(q4.y))), fmaxf(fabsf(q4.z), fabsf(q4.w)));
        const float e4[4] = {q4.x, q4.y, q4.z, q4.w};
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (i + c >= kW2 - (kPre - 1)) ss = fmaf(e4[c], e4[c], ss);
    }
    nxt = locate(pair + gridDim.x);
    fetch(nxt, qn);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        xmax = fmaxf(xmax, __shfl_xor_sync(kFull, xmax, o));
        ss += __shfl_xor_sync(kFull, ss, o);
    }
    if (lane == 0) {
        sm.red[warp] = xmax;
        sm.red2[warp] = ss;
    }
    __syncthreads();
    xmax = 0.0f;
    ss = 0.0f;
#pragma unroll
    for (int w = 0; w < kFT / 32; ++w) {
        xmax = fmaxf(xmax, sm.red[w]);
        ss += sm.red2[w];
    }
    if (!(xmax <= 3.0e38f)) xmax = 3.0e38f;  // NaN / inf samples: everything in the tile pair becomes a candidate
    // pack the two tiles: z[i] = x_a[t0a - 439 + i] + i x_b[t0b - 439 + i]
    for (int i = tid; i < kN; i += kFT)
        sm.buf[0][i] = i < kZ2 ? make_float2(xs0[kW2 - (kPre - 1) + i], xs0[kXs2 + kW2 - (kPre - 1) + i]) : make_float2(0.f, 0.f);

    // ---- EMA power (as in ask_scan_kernel), each half over its own tile
    // a thread's kPer consecutive samples are read ONCE, as 16-byte loads (scalar loads at a 64-byte thread stride are
    // 16-way bank conflicts), squared and scaled, and kept in registers for both passes of the scan
    static_assert(kPer % 4 == 0 && kW2 % kPer == 0, "EMA blocking");
    float e[kPer];
#pragma unroll
    for (int j = 0; j < kPer; j += 4) {
        const float4 v4 = *reinterpret_cast<const float4*>(xs + htid * kPer + j);
        e[j] = v4.x * v4.x * kInv64;
        e[j + 1] = v4.y * v4.y * kInv64;
        e[j + 2] = v4.z * v4.z * kInv64;
        e[j + 3] = v4.w * v4.w * kInv64;
    }
    float q = 0.0f;
#pragma unroll
    for (int j = 0; j < kPer; ++j) q = fmaf(q, kC, e[j]);
    float v = q, mul = mul0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const float t = __shfl_up_sync(kFull, v, o);
        if (lane >= o) v = fmaf(t, mul, v);
        mul *= mul;
    }
    if (lane == 31) sm.carry[half][hwarp] = v;
    __syncthreads();
    float pin = 0.0f;
    for (int w = 0; w < hwarp; ++w) pin = fmaf(pin, c768, sm.carry[half][w]);
    float prev = __shfl_up_sync(kFull, v, 1);
    if (lane == 0) prev = 0.0f;
    float p = fmaf(pin, clane, prev);
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
        p = fmaf(p, kC, e[j]);
        e[j] = p;
    }
    if (htid * kPer >= kW2) {  // all of the thread's samples are positions of the tile (kW2 is a multiple of kPer)
#pragma unroll
        for (int j = 0; j < kPer; j += 4)
            *reinterpret_cast<float4*>(&sm.ps[half][htid * kPer - kW2 + j]) = make_float4(e[j], e[j + 1], e[j + 2], e[j + 3]);
    }
    __syncthreads();

    // ---- fast correlation
    fft2048<false>(sm.buf[0], sm.buf[1], sm.tw, sm.hc, tid);
    fft2048<true>(sm.buf[0], sm.buf[1], sm.tw, nullptr, tid);

    // ---- candidate test; position t0 + m is z index m
    const float tol = kFftRel * h_scale * sqrtf(ss) * 1.001f + 6.5e-5f * xmax + 1e-7f;
#pragma unroll 1
    for (int chunk = htid; chunk < kT2 / 8; chunk += 128) {  // 8 positions per thread; four lanes cover a 32-sample word
        uint32_t any = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int m = 8 * chunk + j;
            const int64_t i = t0 + m;
            const float2 y = sm.buf[0][m];
            const float sp = (half ? y.y : y.x) + tol;
            const float pw = sm.ps[half][m];
            const bool cand = valid && i < n && (!(sp <= 0.05f) && !(sp <= 2.0f * pw * (1.0f - 1e-4f)));  // NaN -> candidate
            any |= cand ? 1u : 0u;
        }
        any |= __shfl_xor_sync(kFull, any, 1);
        any |= __shfl_xor_sync(kFull, any, 2);
        if ((lane & 3) == 0 && any) {
            const uint64_t wbit = sd.word_base + uint64_t(t0 + 8 * chunk) / 32;
            atomicOr(wflag + (wbit >> 5), 1u << (wbit & 31));
        }
    }
    __syncthreads();  // the next pair reuses xs / ps / buf
    }
}

// ---------------------------------------------------------------- exact power chain, in parallel
// `power` (ask.rs:696-697) is one serial chain per capture: p <- rn(rn(p c) + q), c = 63/64, q = rn(rn(x x) / 64) >= 0.
// The step is MONOTONE in p (every rounding is), and it contracts: two chains fed with the same samples, one started
// BELOW the reference's value and one ABOVE it, keep the reference's chain between them and approach each other by 63/64
// per step until they land on the same float — from there on both ARE the reference's chain, bit for bit. So a thread
// owns a segment of kSeg samples, starts kWarm samples before it from 0 and from a value that is not below the reference's
// power there (see ask_power_kernel), and stores the power at every 32-sample
// word boundary of its segment once the two have met (-1 = not known: digital silence, where 0 and the decayed upper
// chain sit on different subnormal fixed points, the first ~1500 samples after it, NaN / inf captures — the walk then
// runs the chain literally from its last exact position, as before). The walk reads the checkpoints instead of chaining
// through every sample of every capture with 32 lanes computing one scalar.
#ifndef TMK_PW_SEG
#define TMK_PW_SEG 8192
#endif
constexpr int kSeg = TMK_PW_SEG, kWarm = 2048;
__device__ __forceinline__ uint32_t seg_stream(const AskStream* __restrict__ streams, uint32_t n_streams, uint32_t seg) {
    uint32_t lo = 0, hi = n_streams - 1;
    while (lo < hi) {
        const uint32_t mid = (lo + hi + 1) >> 1;
        if (streams[mid].seg_base <= seg) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// The upper chain needs a value that is certainly not below the reference's power where it starts. The power is a convex
// combination of the squares of ALL earlier samples (p = sum_j w_j x_j^2, sum w_j <= 1; with roundings: <= (1 + 130 u) max),
// so max x^2 over everything before would do — but that is one more pass over the samples (round 2 first had it:
// ask_xmax_kernel, 1.2 ms at C2, 7.9 GB). Instead the thread looks at the kBound samples in front of its bracket, starts the
// upper chain from kHead x their largest square, and ask_power_check_kernel verifies the assumption afterwards from the
// segments' own maxima (which the threads record as they go): the power at the bracket's start is at most
//   1.0002 (M_{k-1} + c^6144 M_{k-2} + c^14336 max_{j <= k-3} M_j),   M_j = the largest square in segment j, c = 63/64,
// (segment k-1 holds the bracket; c^6144 = 6.2e-43; the last term is below 1e-58 for finite samples) — if that exceeds the
// start value, or any earlier segment holds a NaN / inf (the reference's power is then NaN / inf for good), the segment's
// checkpoints are withdrawn (-1: the walk chains literally there, the filter reads 0). A segment whose surroundings are more
// than sqrt(kHead) = 8 x louder than the 512 samples in front of its bracket is the only one that pays.
constexpr int kBound = 512;        // samples in front of the bracket whose maximum gives the start value of the upper chain
constexpr float kHead = 64.0f;     // head room of that start value (64 x: ~260 more steps until the chains meet)

// Staging: a lane owns a segment, so the lanes of a warp read 32 KB apart — as plain loads that is 32 cache lines per request
// (2.9 ms at C2: bound by the L1 tag rate of fully divergent requests, DRAM at 52 %). Instead every lane moves ITS OWN next
// 512 bytes into shared memory with one bulk copy (cp.async.bulk, completion on the warp's mbarrier of that stage; rows
// 528 bytes apart so that the lanes' 16-byte reads fall on different bank groups) while the warp runs the chains over the
// previous stage: no register staging, no L1, a lane never touches another lane's row.
#ifndef TMK_PW_CHUNK
#define TMK_PW_CHUNK 128
#define TMK_PW_STAGES 2
#define TMK_PW_WARPS 2
#endif
constexpr int kPwChunk = TMK_PW_CHUNK;            // samples per lane and stage
constexpr int kPwPitch = kPwChunk * 4 + 16;       // bytes between the rows of two lanes
constexpr int kPwStages = TMK_PW_STAGES;
constexpr int kPwWarps = TMK_PW_WARPS;
constexpr int kPwRounds = (kBound + kWarm + kSeg) / kPwChunk;
constexpr int kPwSmem = kPwWarps * kPwStages * 32 * kPwPitch;
static_assert((kBound + kWarm) % kPwChunk == 0 && kSeg % kPwChunk == 0, "rounds are whole chunks");

__device__ __forceinline__ uint32_t pw_smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }

__global__ void __launch_bounds__(kPwWarps * 32)
ask_power_kernel(const float* __restrict__ samples, const AskStream* __restrict__ streams, uint32_t n_streams, uint32_t seg0, uint32_t n_segs,
                 float* __restrict__ pchk, uint32_t* __restrict__ seg_max, float* __restrict__ seg_u0, uint32_t* __restrict__ first_bad) {
    extern __shared__ __align__(128) unsigned char pw_rows[];
    __shared__ __align__(8) uint64_t pw_bar[kPwWarps][kPwStages];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        for (int s = 0; s < kPwStages; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pw_smem_u32(&pw_bar[warp][s])), "r"(32) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint32_t seg = seg0 + (blockIdx.x * kPwWarps + warp) * 32 + lane;
    const bool live = seg < n_segs;
    uint32_t si = 0;
    int64_t n = 0, s0 = 0, s1 = 0, e = 0;
    const float* __restrict__ x = samples;
    float* __restrict__ chk = pchk;
    uint32_t seg_in = 0;
    if (live) {
        si = seg_stream(streams, n_streams, seg);
        const AskStream sd = streams[si];
        n = int64_t(sd.n);
        seg_in = seg - sd.seg_base;
        s0 = int64_t(seg_in) * kSeg;
        s1 = min(s0 + int64_t(kSeg), n);
        e = min(s0 + int64_t(kSeg), n & ~int64_t(31));  // end of the whole words: the chain runs over [.., e)
        x = samples + sd.offset;
        chk = pchk + sd.word_base;  // chk[w] = power BEFORE sample 32 w
    }
    const int64_t base = s0 - (kWarm + kBound);  // first sample of round 0
    const bool first = base <= 0;                // the chain's own start: exact from sample 0, nothing in front of it
    unsigned char* const my_rows = pw_rows + size_t(warp) * (kPwStages * 32 * kPwPitch) + size_t(lane) * kPwPitch;
    auto issue = [&](int r) {
        const int st = r % kPwStages;
        const int64_t a = base + int64_t(r) * kPwChunk;
        int64_t cnt = min(int64_t(kPwChunk), e - a);
        if (!live || a < 0 || cnt < 0) cnt = 0;
        const uint32_t bytes = uint32_t(cnt) * 4u, bar = pw_smem_u32(&pw_bar[warp][st]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        if (bytes)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             pw_smem_u32(my_rows + size_t(st) * (32 * kPwPitch))),
                         "l"(x + a), "r"(bytes), "r"(bar)
                         : "memory");
        return int(cnt >> 5);  // whole words of this round
    };
    float pl = 0.0f, pu = 0.0f;
    bool merged = first;
    uint32_t mb = 0;    // largest |x| of the kBound samples in front of the bracket
    uint32_t smax = 0;  // largest |x| of the segment's own samples, as float bits (NaN > inf > finite)
    if (live && first) seg_u0[seg] = -1.0f;  // nothing to verify
    int words[kPwStages];
#pragma unroll
    for (int r = 0; r < kPwStages - 1; ++r) words[r] = issue(r);
#pragma unroll 1
    for (int r0 = 0; r0 < kPwRounds; r0 += kPwStages) {
#pragma unroll
        for (int k = 0; k < kPwStages; ++k) {
            const int r = r0 + k;
            if (r >= kPwRounds) break;
            if (r + kPwStages - 1 < kPwRounds) words[(k + kPwStages - 1) % kPwStages] = issue(r + kPwStages - 1);
            {
                const uint32_t bar = pw_smem_u32(&pw_bar[warp][k]), parity = uint32_t(r0 / kPwStages) & 1u;
                asm volatile(
                    "{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(bar),
                    "r"(parity)
                    : "memory");
            }
            const int64_t a = base + int64_t(r) * kPwChunk;
            const float4* __restrict__ row = reinterpret_cast<const float4*>(my_rows + size_t(k) * (32 * kPwPitch));
            const int nw = words[k];
            const bool both = !__all_sync(0xffffffffu, merged || nw == 0);
            for (int w = 0; w < nw; ++w) {
                const int64_t pos = a + 32 * w;
                float4 v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = row[w * 8 + j];
                if (!first && pos < base + kBound) {  // in front of the bracket: only the maximum
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        mb = max(max(mb, __float_as_uint(v[j].x) & 0x7fffffffu), max(__float_as_uint(v[j].y) & 0x7fffffffu,
                                 max(__float_as_uint(v[j].z) & 0x7fffffffu, __float_as_uint(v[j].w) & 0x7fffffffu)));
                    continue;
                }
                if (!first && pos == base + kBound) {  // the bracket opens
                    const float xm = __uint_as_float(mb);
                    pu = __fmaf_rn(kHead * 1.001f, __fmul_rn(xm, xm), 1e-30f);  // (inf / NaN: the chains never meet)
                    seg_u0[seg] = pu;
                }
                if (!merged) merged = pl == pu;
                if (pos >= s0) {
                    chk[pos >> 5] = merged ? pl : -1.0f;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        smax = max(max(smax, __float_as_uint(v[j].x) & 0x7fffffffu), max(__float_as_uint(v[j].y) & 0x7fffffffu,
                                   max(__float_as_uint(v[j].z) & 0x7fffffffu, __float_as_uint(v[j].w) & 0x7fffffffu)));
                }
                if (!both) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        pl = __fadd_rn(__fmul_rn(pl, kC), __fmul_rn(__fmul_rn(v[j].x, v[j].x), kInv64));
                        pl = __fadd_rn(__fmul_rn(pl, kC), __fmul_rn(__fmul_rn(v[j].y, v[j].y), kInv64));
                        pl = __fadd_rn(__fmul_rn(pl, kC), __fmul_rn(__fmul_rn(v[j].z, v[j].z), kInv64));
                        pl = __fadd_rn(__fmul_rn(pl, kC), __fmul_rn(__fmul_rn(v[j].w, v[j].w), kInv64));
                    }
                    pu = pl;  // (all lanes of the warp have met)
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float q0 = __fmul_rn(__fmul_rn(v[j].x, v[j].x), kInv64), q1 = __fmul_rn(__fmul_rn(v[j].y, v[j].y), kInv64);
                        const float q2 = __fmul_rn(__fmul_rn(v[j].z, v[j].z), kInv64), q3 = __fmul_rn(__fmul_rn(v[j].w, v[j].w), kInv64);
                        pl = __fadd_rn(__fmul_rn(pl, kC), q0);
                        pu = __fadd_rn(__fmul_rn(pu, kC), q0);
                        pl = __fadd_rn(__fmul_rn(pl, kC), q1);
                        pu = __fadd_rn(__fmul_rn(pu, kC), q1);
                        pl = __fadd_rn(__fmul_rn(pl, kC), q2);
                        pu = __fadd_rn(__fmul_rn(pu, kC), q2);
                        pl = __fadd_rn(__fmul_rn(pl, kC), q3);
                        pu = __fadd_rn(__fmul_rn(pu, kC), q3);
                    }
                }
            }
            __syncwarp();
        }
    }
    if (!live) return;
    if (e < s1) {  // the capture's partial last word: its checkpoint, and its samples for the maximum
        if (!merged) merged = pl == pu;
        chk[e >> 5] = merged ? pl : -1.0f;
        for (int64_t i = e; i < s1; ++i) smax = max(smax, __float_as_uint(x[i]) & 0x7fffffffu);
    }
    seg_max[seg] = smax;
    // NaN / inf, or a sample whose square overflows (the reference's power is inf from there on): nothing behind it can be bracketed
    if (smax >= 0x5d5e0b6bu /* 1e18f */) atomicMin(first_bad + si, seg_in);
}

// see above: withdraws the checkpoints of the segments whose upper chain may have started too low
__global__ void __launch_bounds__(128)
ask_power_check_kernel(const AskStream* __restrict__ streams, uint32_t n_streams, uint32_t seg0, uint32_t n_segs,
                       const uint32_t* __restrict__ seg_max, const float* __restrict__ seg_u0, const uint32_t* __restrict__ first_bad,
                       float* __restrict__ pchk) {
    const uint32_t seg = seg0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (seg >= n_segs) return;
    const float u0 = seg_u0[seg];
    if (u0 < 0.0f) return;  // started at the capture's first sample
    const uint32_t si = seg_stream(streams, n_streams, seg);
    const AskStream sd = streams[si];
    const uint32_t k = seg - sd.seg_base;  // >= 1 here
    bool ok = first_bad[si] >= k;          // no NaN / inf / overflowing square in the segments before this one
    if (ok) {
        const double m1 = double(__uint_as_float(seg_max[seg - 1]));
        const double m2 = k >= 2 ? double(__uint_as_float(seg_max[seg - 2])) : 0.0;
        const double bound = 1.0002 * (m1 * m1 + 6.3e-43 * m2 * m2) + 1e-40;  // (older segments: below 1e-58 for finite samples)
        ok = bound <= double(u0);
    }
    if (ok) return;
    const int64_t s0 = int64_t(k) * kSeg, s1 = min(s0 + int64_t(kSeg), int64_t(sd.n));
    float* __restrict__ chk = pchk + sd.word_base;
    for (int64_t w = s0 >> 5; w * 32 < s1; ++w) chk[w] = -1.0f;
}

// ---------------------------------------------------------------- walk
// scratch of the walk kernel at namespace scope: every access compiles to LDS / STS with immediate offsets (through the
// generic pointers of AskCtx the taps of the exact dot were LD.E with 64-bit address arithmetic, ~6 instructions per tap)
__shared__ __align__(16) float g_wpre[kPre];       // the chirp
constexpr int kStageMax = 1024;
__shared__ __align__(16) float g_wxs[4][kStageMax];  // the samples under the windows of a run of words, per warp
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }

struct AskCtx {
    const float* x;
    int64_t n;
    const float* pre;  // shared memory copy
    const float* car;  // global
    int lane;
    // exact EMA chain: `p` is the power after sample cpos - 1 (0 before the first sample)
    int64_t cpos;
    float p;
    float* stage;  // 128 floats of shared memory per warp
    const float* chk;  // exact power before sample 32 w per word of this capture, or -1 (ask_power_kernel); may be null
    int64_t st_lo, st_hi, st_r;  // samples [st_lo, st_hi) are staged in g_wxs (zeros before st_r); st_hi == st_lo: nothing
};

// advance the EMA chain to `target` (power after sample target - 1)            ask.rs:696-697
__device__ __noinline__ void chain_to(AskCtx& c, int64_t target) {
    const float* __restrict__ x = c.x;
    // the nearest checkpoint at or before the target (64 words back at most) replaces the chain up to it
    if (c.chk != nullptr && (target & 31) == 0 && target > c.cpos) {
        const int64_t wt = target >> 5;
#pragma unroll 1
        for (int rnd = 0; rnd < 2; ++rnd) {
            const int64_t w = wt - 32 * rnd - c.lane;
            const float v = (w >= 0 && w * 32 > c.cpos) ? c.chk[w] : -1.0f;
            const uint32_t ok = __ballot_sync(kFull, !(v < 0.0f));  // -1 = unknown; a NaN power is a known value
            if (ok) {
                const int l = __ffs(ok) - 1;
                c.p = __shfl_sync(kFull, v, l);
                c.cpos = (wt - 32 * rnd - l) * 32;
                break;
            }
        }
        if (c.cpos == target) return;
    }
    float p = c.p;
    int64_t pos = c.cpos;
    while (pos < target) {
        const int64_t left = target - pos;
        if (left >= 128) {  // four loads in flight per lane: the chain is one exposed memory round trip per 128 samples
            float v0 = x[pos + c.lane], v1 = x[pos + 32 + c.lane], v2 = x[pos + 64 + c.lane], v3 = x[pos + 96 + c.lane];
            v0 = fmul(fmul(v0, v0), kInv64);
            v1 = fmul(fmul(v1, v1), kInv64);
            v2 = fmul(fmul(v2, v2), kInv64);
            v3 = fmul(fmul(v3, v3), kInv64);
            __syncwarp();
            c.stage[c.lane] = v0;
            c.stage[32 + c.lane] = v1;
            c.stage[64 + c.lane] = v2;
            c.stage[96 + c.lane] = v3;
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const float4 q = *reinterpret_cast<const float4*>(c.stage + 4 * j);
                p = fadd(fmul(p, kC), q.x);
                p = fadd(fmul(p, kC), q.y);
                p = fadd(fmul(p, kC), q.z);
                p = fadd(fmul(p, kC), q.w);
            }
            pos += 128;
            continue;
        }
        const int nstep = left < 32 ? int(left) : 32;
        float v = c.lane < nstep ? x[pos + c.lane] : 0.0f;
        v = fmul(fmul(v, v), kInv64);
        __syncwarp();
        c.stage[c.lane] = v;
        __syncwarp();
        if (nstep == 32) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 q = *reinterpret_cast<const float4*>(c.stage + 4 * j);
                p = fadd(fmul(p, kC), q.x);
                p = fadd(fmul(p, kC), q.y);
                p = fadd(fmul(p, kC), q.z);
                p = fadd(fmul(p, kC), q.w);
            }
        } else {
            for (int j = 0; j < nstep; ++j) p = fadd(fmul(p, kC), c.stage[j]);
        }
        pos += nstep;
    }
    c.p = p;
    c.cpos = target;
}

// exact sync_power and power of the 32 positions of word w (position 32 w + lane); `r` is where the FIFO
// was last zeroed (samples before r read as zeros)                                 ask.rs:701-709
__device__ __noinline__ void eval_word(AskCtx& c, int64_t w, int64_t r, float* sp_out, float* pw_out) {
    const int64_t i = w * 32 + c.lane;
    chain_to(c, w * 32);
    // power after each of the next 32 samples, captured by the lane that owns the position
    {
        float v = i < c.n ? c.x[i] : 0.0f;
        v = fmul(fmul(v, v), kInv64);
        float p = c.p, mine = 0.0f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            p = fadd(fmul(p, kC), __shfl_sync(kFull, v, j));
            if (j == c.lane) mine = p;
        }
        *pw_out = mine;
        // the chain itself only advances over real samples
        const int64_t lim = c.n - w * 32;
        if (lim >= 32) {
            c.p = p;
            c.cpos = w * 32 + 32;
        } else {
            c.p = __shfl_sync(kFull, mine, int(lim > 0 ? lim - 1 : 0));
            c.cpos = c.n;
        }
    }
    // the 471 samples under the 32 windows of the word, staged once (coalesced); samples before r read as the zeroed FIFO
    // (a tap on the zeroed FIFO adds +-0.0 to a sum that starts at +0.0 — no effect — so no tap needs a mask; whole groups
    // of 8 that are zero for every lane are skipped: right after a frame most of the window is still empty)
    const int warp = threadIdx.x >> 5;
    float* __restrict__ xs = g_wxs[warp];
    const int64_t base = w * 32 - (kPre - 1);  // sample index of xs[0]
    const float* __restrict__ x = c.x;
    // (the words right behind a frame come one after the other and share 439 of their 471 samples: the whole run — at
    //  most 14 words, 919 samples — is staged when its first word is looked at)
    if (!(c.st_r == r && base >= c.st_lo && base + (kPre + 31) <= c.st_hi)) {
        int64_t hi = base + (kPre + 31);
        const int64_t pf_last = r + (kPre - 2);  // last position whose window is still partial
        if (w * 32 <= pf_last) hi = max(hi, min((pf_last >> 5) * 32 + 32, base + int64_t(kStageMax)));
        __syncwarp();
        for (int64_t idx = base + c.lane; idx < hi; idx += 32) xs[idx - base] = (idx >= r && idx < c.n) ? __ldg(x + idx) : 0.0f;
        __syncwarp();
        c.st_lo = base;
        c.st_hi = hi;
        c.st_r = r;
    }
    const int64_t zeros = r - base;  // taps below this read zeros for lane 0; lane l reads tap k at xs[(base - st_lo) + l + k]
    const int k0 = zeros > 31 ? int(min(zeros - 31, int64_t(kPre))) & ~7 : 0;
    const float* __restrict__ xl = xs + (base - c.st_lo) + c.lane;
    float sum = 0.0f;
#pragma unroll 5
    for (int k = k0; k < kPre; k += 8) {
        const float4 pa = *reinterpret_cast<const float4*>(g_wpre + k), pb = *reinterpret_cast<const float4*>(g_wpre + k + 4);
        sum = fadd(sum, fmul(xl[k + 0], pa.x));
        sum = fadd(sum, fmul(xl[k + 1], pa.y));
        sum = fadd(sum, fmul(xl[k + 2], pa.z));
        sum = fadd(sum, fmul(xl[k + 3], pa.w));
        sum = fadd(sum, fmul(xl[k + 4], pb.x));
        sum = fadd(sum, fmul(xl[k + 5], pb.y));
        sum = fadd(sum, fmul(xl[k + 6], pb.z));
        sum = fadd(sum, fmul(xl[k + 7], pb.w));
    }
    *sp_out = __fdiv_rn(sum, 200.0f);
}

// carrier mixing, 11-tap moving average, integrate and dump of the 108 bits      ask.rs:733-761
__device__ __noinline__ void decode_frame(const AskCtx& c, int64_t s, uint8_t* bits14, uint32_t* id_out) {
    uint32_t words[4];
    const float* __restrict__ x = c.x + s + 1;
    const float* __restrict__ car = c.car;
    for (int rnd = 0; rnd < 4; ++rnd) {
        const int j = rnd * 32 + c.lane;
        bool bit = false;
        if (j < kBits) {
            // the 30 products under the bit's 20 averaging windows, each rounded once (the reference forms every product
            // anew in each of the 11 windows that hold it: same value), then the 11-term sums in the reference's order
            float pr[30];
#pragma unroll
            for (int t = 0; t < 30; ++t) {
                const int k = 5 + j * TMK_ASK_SAMPLES_PER_BIT + t;
                // decode_fifo[k]: rx[s+1 .. s+201) then rx[s+202 ..)  (the sample at s+201 is dropped)
                pr[t] = fmul(__ldg(x + k + (k >= 200 ? 1 : 0)), __ldg(car + k));
            }
            float tot = 0.0f;
#pragma unroll
            for (int t = 0; t < 20; ++t) {
                float sm = 0.0f;
#pragma unroll
                for (int u = 0; u < 11; ++u) sm = fadd(sm, pr[t + u]);
                tot = fadd(tot, __fdiv_rn(sm, 11.0f));
            }
            bit = tot > 0.0f;
        }
        words[rnd] = __brev(__ballot_sync(kFull, bit));  // bit j at the MSB side
    }
    if (c.lane < 14) bits14[c.lane] = uint8_t(words[c.lane >> 2] >> (24 - 8 * (c.lane & 3)));
    *id_out = words[0] >> 24;
}

#ifndef TMK_ASK_WALK_CTAS
#define TMK_ASK_WALK_CTAS 5  // measured at C2 (walk, ms): 3 -> 6.82, 4 -> 6.80, 5 -> 6.38, 6 -> 6.50 (spills), 8 -> 6.88
#endif
__global__ void __launch_bounds__(128, TMK_ASK_WALK_CTAS)
ask_walk_kernel(const float* __restrict__ samples, const AskStream* __restrict__ streams, uint32_t item0, uint32_t n_streams,
                const float* __restrict__ tab, const uint32_t* __restrict__ wflag, const float* __restrict__ pchk,
                tmk_ask_rec* __restrict__ recs, uint32_t* __restrict__ rec_ord, uint32_t* __restrict__ stream_cnt, uint64_t rec_cap,
                unsigned long long* __restrict__ counters, unsigned long long* __restrict__ work_ctr, TmkSeg seg,
                uint32_t* __restrict__ rec_item, uint32_t want_bits) {
    // want_bits == 0: the caller only needs the synchroniser (mcm.cu demodulates the bodies itself): no BPSK demodulation
    // work items [item0, n_streams): a chunk of the batch (the chunks of a batch run on different CUDA streams, so that
    // the power chain / filter / walk of neighbouring chunks overlap); work_ctr: this launch's own zeroed counter
    // seg.items != nullptr: the work items are SEGMENTS of captures (the segmented walk of tmk_internal.cuh, for this
    // receiver: its state right after a frame is the position r where the FIFO was zeroed — power is a function of the
    // samples alone — so a walker that starts mid-capture as "searching, nothing pending" is the true receiver from the
    // first frame end it shares with the walker coming from the segments before it)
    const bool segm = seg.items != nullptr;
    __shared__ __align__(16) float stage_s[4][128];
    for (int i = threadIdx.x; i < kPre; i += blockDim.x) g_wpre[i] = tab[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (;;) {
        uint32_t sidx = 0;
        if (lane == 0) sidx = item0 + uint32_t(atomicAdd(work_ctr, 1ull));
        sidx = __shfl_sync(kFull, sidx, 0);
        if (sidx >= n_streams) return;
        uint32_t real = sidx, ord0 = 0;
        int64_t q0 = 0, q1 = int64_t(0x7fffffffffffffffll);
        bool resumed = false;
        if (segm) {
            const TmkSegItem it = seg.items[sidx];
            real = it.stream;
            if (seg.pass == 1) {
                q0 = int64_t(it.q0);
                q1 = it.q1 == ~0ull ? q1 : int64_t(it.q1);
            } else {
                const TmkSegState* ps = seg.st + sidx;
                if (ps->state != 1) continue;  // only walkers that paused at the end of their segment go on
                q0 = int64_t(ps->q);
                ord0 = ps->ord;
                resumed = true;
            }
        }
        const AskStream sd = streams[real];
        AskCtx c;
        c.x = samples + sd.offset;
        c.n = int64_t(sd.n);
        c.pre = g_wpre;
        c.car = tab + kPre;
        c.lane = lane;
        c.cpos = 0;
        c.p = 0.0f;
        c.stage = stage_s[warp];
        c.chk = pchk != nullptr ? pchk + sd.word_base : nullptr;
        c.st_lo = c.st_hi = c.st_r = 0;
        const int64_t nwords = (c.n + 31) / 32;
        // a walker that resumes (pass 2) stands right behind a frame: FIFO zeroed at q0; one that starts a segment searches
        // with a full FIFO (r = 0) from q0 on
        int64_t r = resumed ? q0 : 0, cur = q0, start = 0;
        float local_max = 0.0f;
        bool done = false;
        uint32_t n_emitted = ord0;  // records of this walker, in time order: its position in the ordered output
        uint32_t seg_exit = 0, sync_item = 0, sync_ord = 0;
        bool at_entry = resumed;    // standing at a frame end (a state the position alone determines)
        while (!done) {
            if (segm && at_entry) {
                at_entry = false;
                if (seg.pass == 1) {
                    if (n_emitted >= 1 && n_emitted <= TMK_SEG_HEADS && lane == 0) seg.st[sidx].head[n_emitted - 1] = uint64_t(r);
                    if (r >= q1) {
                        seg_exit = 1;
                        break;
                    }
                } else {
                    const TmkSegItem it = seg.items[sidx];
                    uint64_t k = uint64_t(r) / it.seg_len;
                    if (k >= it.n_seg) k = it.n_seg - 1;
                    const uint32_t owner = it.first_item + uint32_t(k);
                    const TmkSegState* os = seg.st + owner;
                    const bool hit = uint32_t(lane) < os->n_head && os->head[lane] == uint64_t(r);
                    const uint32_t hm = __ballot_sync(kFull, hit);
                    if (hm && owner != sidx) {
                        sync_item = owner;
                        sync_ord = uint32_t(__ffs(hm));  // head[i] is the frame end behind the owner's record i: valid from i + 1
                        seg_exit = 2;
                        break;
                    }
                }
            }
            // ---- next word to evaluate: everything inside the first 439 positions after the FIFO was zeroed
            //      (partial windows are not covered by the filter), afterwards the flagged words only
            int64_t w = cur >> 5;
            if (cur >= r + (kPre - 1)) {
                int64_t found = -1;
                while (w < nwords && found < 0) {
                    // 32 flag words = 1024 sample words per round
                    const uint64_t b0 = sd.word_base + uint64_t(w);
                    const uint64_t fw0 = b0 >> 5;
                    const uint64_t lastbit = sd.word_base + uint64_t(nwords) - 1;
                    uint32_t fl = 0;
                    if (fw0 + lane <= (lastbit >> 5)) fl = wflag[fw0 + lane];
                    if (lane == 0) fl &= ~0u << (b0 & 31);  // bits before w
                    if (fw0 + lane == (lastbit >> 5) && (lastbit & 31) != 31) fl &= (2u << (lastbit & 31)) - 1u;  // next stream's bits
                    const uint32_t m = __ballot_sync(kFull, fl != 0);
                    if (m) {
                        const int l = __ffs(m) - 1;
                        const uint32_t f = __shfl_sync(kFull, fl, l);
                        found = int64_t(((fw0 + l) << 5) + (__ffs(f) - 1) - sd.word_base);
                    } else {
                        w = int64_t(((fw0 + 32) << 5) - sd.word_base);
                    }
                }
                w = found < 0 ? nwords : found;
            }
            // ---- a pending lock that comes before the next evaluated word            ask.rs:717-723
            int64_t lock = -1;
            if (start != 0 && start + 201 < (w < nwords ? w * 32 : c.n)) lock = start + 201;
            if (lock < 0) {
                if (w >= nwords) break;
                float sp, pw;
                eval_word(c, w, r, &sp, &pw);
                const int64_t i = w * 32 + lane;
                const bool cond0 = i >= cur && i < c.n && sp > fmul(pw, 2.0f) && sp > 0.05f;
                uint32_t m = __ballot_sync(kFull, cond0);
                const int64_t wend = (w * 32 + 32 < c.n) ? w * 32 + 32 : c.n;
                while (m) {
                    const int l = __ffs(m) - 1;
                    m &= m - 1;
                    const int64_t il = w * 32 + l;
                    if (start != 0 && start + 201 < il) { lock = start + 201; break; }
                    const float spl = __shfl_sync(kFull, sp, l);
                    if (spl > local_max) {
                        local_max = spl;
                        start = il;
                    } else if (start != 0 && il > start + 200) {
                        lock = il;
                        break;
                    }
                }
                if (lock < 0 && start != 0 && start + 201 < wend) lock = start + 201;
                if (lock < 0) {
                    cur = w * 32 + 32;
                    if (cur >= c.n) break;
                    continue;
                }
            }
            // ---- locked at `lock` (= start + 201): gather 4752 samples, decode     ask.rs:725-793
            const int64_t s = start;
            if (s + 2 + kFrame > c.n) break;  // the capture ends while the frame is being gathered
            uint32_t id = 0;
            unsigned long long slot = 0;
            if (lane == 0) slot = atomicAdd(&counters[0], 1ull);
            slot = __shfl_sync(kFull, slot, 0);
            __shared__ uint8_t bits_s[4][16];
            if (want_bits) {
                decode_frame(c, s, bits_s[warp], &id);
            } else if (lane < 14) {
                bits_s[warp][lane] = 0;
            }
            __syncwarp();
            if (lane == 0 && slot < rec_cap) {
                tmk_ask_rec rec;
                memset(&rec, 0, sizeof rec);
                rec.stream = real;
                rec.accepted = (id > 0 && id <= 100) ? 1u : 0u;
                rec.start_index = uint64_t(s);
                rec.sync_power = local_max;
                rec.id = uint8_t(id);
                for (int b = 0; b < 14; ++b) rec.bits[b] = bits_s[warp][b];
                recs[slot] = rec;
                rec_ord[slot] = n_emitted;
                if (rec_item != nullptr) rec_item[slot] = sidx;
            }
            ++n_emitted;
            __syncwarp();
            r = s + 2 + kFrame;  // first sample examined in state 0 again
            cur = r;
            start = 0;
            local_max = 0.0f;
            at_entry = true;
            if (cur >= c.n) done = true;
        }
        if (lane == 0) {
            stream_cnt[sidx] = n_emitted;
            if (segm) {
                TmkSegState* ps = seg.st + sidx;
                ps->q = uint64_t(r);
                ps->ord = n_emitted;
                ps->state = seg_exit;
                ps->sync_item = sync_item;
                ps->sync_ord = sync_ord;
                if (seg.pass == 1) ps->n_head = n_emitted < TMK_SEG_HEADS ? n_emitted : TMK_SEG_HEADS;
            }
        }
    }
}

// the ordered scatter of a segmented walk: only the records of walkers on the chain, from their first valid one on
__global__ void ask_seg_order_kernel(const tmk_ask_rec* __restrict__ in, const uint32_t* __restrict__ rec_ord,
                                     const uint32_t* __restrict__ rec_item, uint64_t n, const uint64_t* __restrict__ base,
                                     const uint32_t* __restrict__ first_valid, tmk_ask_rec* __restrict__ out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t item = rec_item[i], f = first_valid[item], o = rec_ord[i];
    if (f == 0xffffffffu || o < f) return;
    out[base[item] + (o - f)] = in[i];
}

// records in (capture, time) order: the walk's warps emit through an atomic slot counter; base = exclusive scan of the
// captures' record counts
__global__ void ask_order_kernel(const tmk_ask_rec* __restrict__ in, const uint32_t* __restrict__ rec_ord, uint64_t n,
                                 const uint64_t* __restrict__ base, tmk_ask_rec* __restrict__ out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const tmk_ask_rec r = in[i];
    out[base[r.stream] + rec_ord[i]] = r;
}

void host_tables(std::vector<float>& tab, size_t n_carrier) {
    // examples/ask.rs:574-605, f32 operation order
    const float pi = 3.14159265358979323846f;
    tab.assign(kPre + n_carrier, 0.0f);
    float fp[kPre];
    for (int i = 0; i < 220; ++i) fp[i] = 2000.0f + (8000.0f * float(i) / 219.0f);
    for (int i = 0; i < 220; ++i) fp[220 + i] = 10000.0f - (8000.0f * float(i) / 219.0f);
    float omega = 0.0f;
    tab[0] = sinf(2.0f * pi * fp[0] * (0.0f / 48000.0f));
    for (int i = 1; i < kPre; ++i) {
        const float dt = float(i) / 48000.0f - float(i - 1) / 48000.0f;
        omega += pi * (fp[i] + fp[i - 1]) * dt;
        tab[i] = sinf(omega);
    }
    for (size_t k = 0; k < n_carrier; ++k) tab[kPre + k] = sinf(2.0f * pi * 10000.0f * (float(k) / 48000.0f));
}

// layout of d_ask_tab: [kPre preamble | kFrame carrier | pad to 8 bytes | kN twiddles (float2) | kN conj(H) / (N 200) (float2)]
constexpr size_t kTabFft = (size_t(kPre) + kFrame + 1) / 2 * 2;  // in floats; twiddles in the per-pass layout of FftSmem::tw

int ensure_tables(tmk_engine* e) {
    if (e->ask_ready) return TMK_OK;
    std::vector<float> tab;
    host_tables(tab, kFrame);
    tab.resize(kTabFft + 4 * kN, 0.0f);
    const double two_pi = 6.283185307179586476925286766559;
    double hmax = 0.0;
    auto put_tw = [&](int slot, double turns) {  // exp(-2 pi i turns)
        tab[kTabFft + 2 * slot] = float(std::cos(two_pi * turns));
        tab[kTabFft + 2 * slot + 1] = float(-std::sin(two_pi * turns));
    };
    for (int p = 1; p <= 256; p *= 4)
        for (int k = 0; k < p; ++k)
            for (int q = 1; q <= 3; ++q) put_tw(p - 1 + (q - 1) * p + k, double(q * k) / double(4 * p));
    for (int k = 0; k < kN / 2; ++k) put_tw(1023 + k, double(k) / double(kN));
    put_tw(kN - 1, 0.0);
    for (int m = 0; m < kN; ++m) {
        double re = 0.0, im = 0.0;  // H[m] = sum_k h[k] exp(-2 pi i m k / N)
        for (int k = 0; k < kPre; ++k) {
            const double a = two_pi * double((int64_t(m) * k) % kN) / kN;
            re += double(tab[k]) * std::cos(a);
            im -= double(tab[k]) * std::sin(a);
        }
        hmax = std::max(hmax, std::sqrt(re * re + im * im));
        tab[kTabFft + 2 * kN + 2 * m] = float(re / (double(kN) * 200.0));
        tab[kTabFft + 2 * kN + 2 * m + 1] = float(-im / (double(kN) * 200.0));  // conj
    }
    e->ask_hscale = float(hmax / 200.0 * 1.0001);
    CK(e->d_ask_tab.reserve(tab.size() * sizeof(float)));
    CK(cudaMemcpyAsync(e->d_ask_tab.p, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    e->ask_ready = true;
    return TMK_OK;
}

}  // namespace

extern "C" {

int tmk_ask_tables(tmk_engine* e, float* preamble440, float* carrier, uint64_t n_carrier) {
    if (!e || !preamble440 || (!carrier && n_carrier)) return tmk_fail(e, TMK_E_INVALID, "tmk_ask_tables: null argument");
    std::vector<float> tab;
    host_tables(tab, size_t(n_carrier));
    std::memcpy(preamble440, tab.data(), kPre * sizeof(float));
    if (n_carrier) std::memcpy(carrier, tab.data() + kPre, size_t(n_carrier) * sizeof(float));
    return TMK_OK;
}

uint64_t tmk_ask_modulated_len(uint32_t n_frames, const uint32_t* gap1, const uint32_t* gap2) {
    uint64_t total = 0;
    for (uint32_t f = 0; f < n_frames; ++f) total += uint64_t(gap1 ? gap1[f] : 0) + kPre + kFrame + (gap2 ? gap2[f] : 0);
    return total;
}

int tmk_ask_modulate(tmk_engine* e, const uint8_t* bits, uint32_t n_frames, const uint32_t* gap1, const uint32_t* gap2,
                     float* out_samples, uint64_t out_cap, uint64_t* out_len) {
    if (!e || (!bits && n_frames) || !out_len || ((!gap1 || !gap2) && n_frames))
        return tmk_fail(e, TMK_E_INVALID, "tmk_ask_modulate: null argument");
    CK(cudaSetDevice(e->device));
    std::vector<uint64_t> off(n_frames + 1, 0);
    for (uint32_t f = 0; f < n_frames; ++f) off[f + 1] = off[f] + gap1[f] + kPre + kFrame + gap2[f];
    const uint64_t total = off[n_frames];
    *out_len = total;
    if (total > out_cap) return tmk_fail(e, TMK_E_CAPACITY, "tmk_ask_modulate: output buffer too small");
    if (total == 0) return TMK_OK;
    if (!out_samples) return tmk_fail(e, TMK_E_INVALID, "tmk_ask_modulate: null output");
    int rc = ensure_tables(e);
    if (rc != TMK_OK) return rc;
    const size_t ob = sizeof(uint64_t) * (n_frames + 1), gb = sizeof(uint32_t) * n_frames, bb = size_t(n_frames) * TMK_ASK_DATA_BITS;
    CK(e->d_misc.reserve(ob + 2 * gb + bb + 64));
    char* dm = static_cast<char*>(e->d_misc.p);
    CK(cudaMemcpyAsync(dm, off.data(), ob, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(dm + ob, gap1, gb, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(dm + ob + gb, gap2, gb, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(dm + ob + 2 * gb, bits, bb, cudaMemcpyHostToDevice, e->stream));
    float* d_out = out_samples;
    const bool dev_out = tmk_is_device_ptr(out_samples);
    if (!dev_out) {
        CK(e->d_samples.reserve(total * sizeof(float)));
        d_out = static_cast<float*>(e->d_samples.p);
    }
    ask_modulate_kernel<<<n_frames, 256, 0, e->stream>>>(static_cast<const float*>(e->d_ask_tab.p),
                                                         reinterpret_cast<const uint8_t*>(dm + ob + 2 * gb),
                                                         reinterpret_cast<const uint64_t*>(dm),
                                                         reinterpret_cast<const uint32_t*>(dm + ob),
                                                         reinterpret_cast<const uint32_t*>(dm + ob + gb), d_out);
    e->launches += 1;
    CK(cudaGetLastError());
    if (!dev_out) CK(cudaMemcpyAsync(out_samples, d_out, total * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

}  // extern "C"

// tables of the chirp + BPSK modem on the device (e->d_ask_tab: [440 chirp | 4752 carrier | FFT tables])
int tmk_ask_ensure_tables(tmk_engine* e) { return ensure_tables(e); }

// The synchroniser + BPSK demodulator on the device: records are left in e->d_ask_recs (unsorted: the warps emit through
// an atomic slot counter), *dx_out is where the samples live on the device. Shared with mcm.cu, whose frames use the
// same chirp preamble and frame envelope.
int tmk_ask_sync_device(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams, uint32_t n_streams,
                        const float** dx_out, uint64_t* n_recs, uint64_t* dev_cap_out, bool want_bits) {
    *n_recs = 0;
    CK(cudaSetDevice(e->device));
    int rc = ensure_tables(e);
    if (rc != TMK_OK) return rc;
    // TMK_ASK_DIRECT=1 keeps the direct-form filter (A/B and cross-check of the FFT form)
    static const bool direct = getenv("TMK_ASK_DIRECT") != nullptr;
    const uint64_t tile_len = direct ? kT : kT2;
    std::vector<AskStream> sd(n_streams);
    uint64_t words = 0, tiles = 0, segs = 0;
    for (uint32_t s = 0; s < n_streams; ++s) {
        if (streams[s].offset % 4 != 0) return tmk_fail(e, TMK_E_INVALID, "stream offset must be a multiple of 4 samples");
        if (streams[s].offset > total_samples || streams[s].n_samples > total_samples - streams[s].offset)
            return tmk_fail(e, TMK_E_INVALID, "stream " + std::to_string(s) + " exceeds the sample buffer");
        sd[s].offset = streams[s].offset;
        sd[s].n = streams[s].n_samples;
        sd[s].word_base = words;
        sd[s].tile_base = uint32_t(tiles);
        sd[s].seg_base = uint32_t(segs);
        segs += (streams[s].n_samples + kSeg - 1) / kSeg;
        if (segs > 0x7fffffffull) return tmk_fail(e, TMK_E_INVALID, "batch too large (segment count)");
        words += (streams[s].n_samples + 31) / 32;
        tiles += (streams[s].n_samples + tile_len - 1) / tile_len;
        if (tiles > 0x7fffffffull) return tmk_fail(e, TMK_E_INVALID, "batch too large (tile count)");
    }
    const float* dx = samples;
    if (total_samples && !tmk_is_device_ptr(samples)) {
        CK(e->d_samples.reserve(total_samples * sizeof(float)));
        CK(cudaMemcpyAsync(e->d_samples.p, samples, total_samples * sizeof(float), cudaMemcpyHostToDevice, e->stream));
        dx = static_cast<const float*>(e->d_samples.p);
    }
    if (reinterpret_cast<uintptr_t>(dx) % 16 != 0) return tmk_fail(e, TMK_E_INVALID, "sample buffer must be 16-byte aligned");
    const size_t flag_words = size_t(words / 32 + 2);
    CK(e->d_streams.reserve(sizeof(AskStream) * n_streams));
    CK(e->d_ask_bits.reserve(flag_words * sizeof(uint32_t)));
    CK(e->d_ask_l2.reserve(16 * sizeof(unsigned long long)));
    uint64_t dev_cap = std::max<uint64_t>(1024, total_samples / 4096 + 2ull * n_streams);
    CK(e->d_ask_recs.reserve(dev_cap * sizeof(tmk_ask_rec)));
    // unordered records + their ordinals | per-capture counts and bases (the ordered records land in d_ask_recs)
    CK(e->d_ask_recs2.reserve(dev_cap * (sizeof(tmk_ask_rec) + sizeof(uint32_t)) + 64 + sizeof(uint32_t) * n_streams + sizeof(uint64_t) * (n_streams + 1) + 64));
    tmk_ask_rec* raw_recs = static_cast<tmk_ask_rec*>(e->d_ask_recs2.p);
    uint32_t* raw_ord = reinterpret_cast<uint32_t*>(raw_recs + dev_cap);
    uint64_t* rec_base = reinterpret_cast<uint64_t*>(static_cast<char*>(e->d_ask_recs2.p) + (dev_cap * (sizeof(tmk_ask_rec) + sizeof(uint32_t)) + 63) / 64 * 64);
    uint32_t* rec_cnt = reinterpret_cast<uint32_t*>(rec_base + n_streams + 1);
    CK(cudaMemcpyAsync(e->d_streams.p, sd.data(), sizeof(AskStream) * n_streams, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemsetAsync(e->d_ask_bits.p, 0, flag_words * sizeof(uint32_t), e->stream));
    CK(cudaMemsetAsync(e->d_ask_l2.p, 0, 16 * sizeof(unsigned long long), e->stream));
    const AskStream* ds = static_cast<const AskStream*>(e->d_streams.p);
    const float* tab = static_cast<const float*>(e->d_ask_tab.p);
    tmk_flush_stage_ms(e);
    CK(cudaEventRecord(e->ev[0], e->stream));
    // exact power checkpoints for the walk (TMK_ASK_NO_CHK=1: the walk chains through every sample itself, for A/B)
    static const bool no_chk = getenv("TMK_ASK_NO_CHK") != nullptr;
    // the filter: tensor-core Toeplitz GEMM (ask_tc.cu) | TMK_ASK_FFT=1: FFT fast correlation | TMK_ASK_DIRECT=1: direct form
    static const bool use_fft = getenv("TMK_ASK_FFT") != nullptr;
    float* pchk = nullptr;
    uint32_t *seg_max = nullptr, *first_bad = nullptr;
    float* seg_u0 = nullptr;
    if (!no_chk && segs) {
        CK(e->d_ask_chk.reserve((words + 2) * sizeof(float) + sizeof(uint32_t) * (2 * segs + n_streams)));
        pchk = static_cast<float*>(e->d_ask_chk.p);
        seg_max = reinterpret_cast<uint32_t*>(pchk + words + 2);
        seg_u0 = reinterpret_cast<float*>(seg_max + segs);
        first_bad = reinterpret_cast<uint32_t*>(seg_u0 + segs);
        CK(cudaMemsetAsync(first_bad, 0xff, sizeof(uint32_t) * n_streams, e->stream));
    }
    const bool tc = pchk != nullptr && !use_fft && !direct;
    std::vector<uint32_t> tc_base(size_t(n_streams) + 1, 0);
    if (tc) {
        std::vector<uint64_t> so(n_streams), sn(n_streams), sw(n_streams);
        for (uint32_t s = 0; s < n_streams; ++s) {
            so[s] = sd[s].offset;
            sn[s] = sd[s].n;
            sw[s] = sd[s].word_base;
        }
        rc = tmk_ask_tc_prepare(e, so.data(), sn.data(), sw.data(), n_streams, tc_base.data());
        if (rc != TMK_OK) return rc;
    }
    static int walk_ctas = 0;
    if (walk_ctas == 0) {
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ask_walk_kernel, 128, 0);
        walk_ctas = sms * (per_sm > 0 ? per_sm : 1);
    }
    // Segmented walk: only the captures that lock have work, and a capture is one serial chain of frames — with few or long
    // captures half of the SMs' warp slots stay empty (C2: 14 busy warps per SM). Captures are cut into segments walked at
    // once and stitched exactly (tmk_internal.cuh; this receiver's state behind a frame is its position).
    std::vector<TmkSegItem> items;
    std::vector<uint32_t> item_of(size_t(n_streams) + 1, 0);  // first walker of every capture
    {
        uint64_t target = 148ull * 28 * 3, seg_min = 65536;  // measured at C2 (walk ms): 6216 walkers 8.1, 12 500 7.1, 25 000 6.7
        if (const char* ev = getenv("TMK_ASK_SEGMENTS")) target = strtoull(ev, nullptr, 10);  // 0 / 1: off
        if (const char* ev = getenv("TMK_ASK_SEG_MIN")) seg_min = std::max<uint64_t>(8192, strtoull(ev, nullptr, 10));
        // (only while the captures alone cannot fill the warp slots — 148 SMs x 28 warps: at 8192 busy captures the
        //  duplicated frames and the second pass cost more than the segments gain: C3 FSK walk 51.8 -> 59.0 ms)
        if (total_samples && uint64_t(n_streams) * 3 <= target && pchk != nullptr) {
            for (uint32_t s = 0; s < n_streams; ++s) {
                const uint64_t n = sd[s].n;
                const uint64_t share = (target * n + total_samples - 1) / total_samples;
                uint64_t nseg = std::max<uint64_t>(1, std::min<uint64_t>(std::min<uint64_t>(share, n / seg_min), 60000));
                const uint64_t seg_len = nseg > 1 ? ((n + nseg - 1) / nseg + 31) / 32 * 32 : std::max<uint64_t>(n, 1);
                nseg = nseg > 1 ? (n + seg_len - 1) / seg_len : 1;
                const uint32_t first = uint32_t(items.size());
                item_of[s] = first;
                for (uint64_t k = 0; k < nseg; ++k) {
                    TmkSegItem it;
                    it.stream = s;
                    it.seg = uint32_t(k);
                    it.first_item = first;
                    it.n_seg = uint32_t(nseg);
                    it.q0 = k * seg_len;
                    it.q1 = k + 1 < nseg ? (k + 1) * seg_len : ~0ull;
                    it.seg_len = seg_len;
                    items.push_back(it);
                }
            }
            item_of[n_streams] = uint32_t(items.size());
            if (items.size() == n_streams) items.clear();
        }
    }
    const bool seg_mode = !items.empty();
    const uint32_t n_items = seg_mode ? uint32_t(items.size()) : n_streams;
    TmkSeg seg{nullptr, nullptr, 0, 0};
    uint32_t *raw_item = nullptr, *first_valid = nullptr, *n_valid = nullptr;
    if (seg_mode) {
        const size_t ib = sizeof(TmkSegItem) * n_items, sb = sizeof(TmkSegState) * n_items;
        const size_t o_item = (ib + sb + 63) / 64 * 64, o_fv = o_item + (sizeof(uint32_t) * dev_cap + 63) / 64 * 64;
        const size_t o_nv = o_fv + (sizeof(uint32_t) * n_items + 63) / 64 * 64, o_cnt = o_nv + (sizeof(uint32_t) * n_items + 63) / 64 * 64;
        const size_t o_base = o_cnt + (sizeof(uint32_t) * n_items + 63) / 64 * 64;
        CK(e->d_ask_seg.reserve(o_base + sizeof(uint64_t) * (n_items + 1) + 64));
        char* b = static_cast<char*>(e->d_ask_seg.p);
        seg.items = reinterpret_cast<const TmkSegItem*>(b);
        seg.st = reinterpret_cast<TmkSegState*>(b + ib);
        raw_item = reinterpret_cast<uint32_t*>(b + o_item);
        first_valid = reinterpret_cast<uint32_t*>(b + o_fv);
        n_valid = reinterpret_cast<uint32_t*>(b + o_nv);
        rec_cnt = reinterpret_cast<uint32_t*>(b + o_cnt);   // per walker
        rec_base = reinterpret_cast<uint64_t*>(b + o_base);
        CK(cudaMemcpyAsync(b, items.data(), ib, cudaMemcpyHostToDevice, e->stream));
        CK(cudaMemsetAsync(b + ib, 0, sb, e->stream));
    }
    if (!seg_mode)
        for (uint32_t s = 0; s <= n_streams; ++s) item_of[s] = s;
    // Chunks of whole captures, each on its own CUDA stream (xmax -> power chain -> filter -> walk of a chunk in order,
    // neighbouring chunks free to overlap), only on request (TMK_ASK_CHUNKS): MEASURED AND NOT SHIPPED — the stages load
    // different units (DRAM, DRAM latency, tensor pipe, issue slots of a few busy warps), but the filter's 227 KB of shared
    // memory per SM keeps it from sharing an SM with the walk's CTAs, and a chunk's walk has fewer captures to fill the
    // warp slots with: C2 at 1 / 2 / 4 chunks 128 / 129 / 114 G samples/s, C3 FSK walk 59 -> 72 ms at 4 chunks.
    constexpr uint32_t kMaxChunks = 4;
    uint32_t n_chunks = 1;
    if (tc) {
        if (const char* ev = getenv("TMK_ASK_CHUNKS")) n_chunks = uint32_t(std::min<uint64_t>(kMaxChunks, std::max<uint64_t>(1, strtoull(ev, nullptr, 10))));
        n_chunks = std::min(n_chunks, n_streams);
    }
    uint32_t cut[kMaxChunks + 1];
    cut[0] = 0;
    {
        uint64_t acc = 0;
        uint32_t g = 1;
        for (uint32_t s = 0; s < n_streams && g < n_chunks; ++s) {
            acc += sd[s].n;
            if (acc * n_chunks >= uint64_t(g) * total_samples && s + 1 < n_streams) cut[g++] = s + 1;
        }
        while (g <= n_chunks) cut[g++] = n_streams;
    }
    unsigned long long* ctr = static_cast<unsigned long long*>(e->d_ask_l2.p);  // [0] record slots, [2 + 2 g + pass - 1] work counters
    if (n_chunks > 1) CK(cudaEventRecord(e->cev[0], e->stream));
    for (uint32_t g = 0; g < n_chunks; ++g) {
        const uint32_t s0 = cut[g], s1 = cut[g + 1];
        if (s1 <= s0) continue;
        cudaStream_t st = n_chunks > 1 ? e->walk_streams[g % 4] : e->stream;
        if (n_chunks > 1) CK(cudaStreamWaitEvent(st, e->cev[0], 0));
        if (pchk != nullptr) {
            const uint32_t g0 = sd[s0].seg_base, g1 = s1 < n_streams ? sd[s1].seg_base : uint32_t(segs);
            if (g1 > g0) {
                static TmkPerDevice pw_attr;
                if (!pw_attr.done(e->device)) {
                    CK(cudaFuncSetAttribute(ask_power_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPwSmem));
                    pw_attr.set(e->device);
                }
                ask_power_kernel<<<(g1 - g0 + kPwWarps * 32 - 1) / (kPwWarps * 32), kPwWarps * 32, kPwSmem, st>>>(dx, ds, n_streams, g0, g1, pchk, seg_max, seg_u0, first_bad);
                ask_power_check_kernel<<<(g1 - g0 + 127) / 128, 128, 0, st>>>(ds, n_streams, g0, g1, seg_max, seg_u0, first_bad, pchk);
                e->launches += 2;
            }
        }
        if (tc) {
            rc = tmk_ask_tc_filter(e, dx, total_samples, n_streams, tc_base[s0], tc_base[s1], tab, pchk,
                                   static_cast<uint32_t*>(e->d_ask_bits.p), st);
            if (rc != TMK_OK) return rc;
        } else if (tiles) {
            if (direct) {
                ask_scan_kernel<<<uint32_t(tiles), kThreads, 0, st>>>(dx, ds, n_streams, tab, static_cast<uint32_t*>(e->d_ask_bits.p));
            } else {
                static TmkPerDevice fft_attr;
                if (!fft_attr.done(e->device)) {
                    CK(cudaFuncSetAttribute(ask_scan_fft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(FftSmem))));
                    fft_attr.set(e->device);
                }
                const float2* twg = reinterpret_cast<const float2*>(tab + kTabFft);
                static int fft_ctas = 0;
                if (fft_ctas == 0) {
                    int sms = 0, per_sm = 0;
                    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device);
                    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ask_scan_fft_kernel, kFT, sizeof(FftSmem));
                    fft_ctas = sms * (per_sm > 0 ? per_sm : 1);
                }
                ask_scan_fft_kernel<<<uint32_t(std::min<uint64_t>((tiles + 1) / 2, uint64_t(fft_ctas))), kFT, sizeof(FftSmem), st>>>(
                    dx, ds, n_streams, uint32_t(tiles), twg, twg + kN, e->ask_hscale, static_cast<uint32_t*>(e->d_ask_bits.p));
            }
            e->launches += 1;
        }
        CK(cudaGetLastError());
        if (g + 1 == n_chunks) CK(cudaEventRecord(e->ev[1], st));  // the last filter is done: "filter" / "walk" split of the stage times
        const uint32_t i0 = item_of[s0], i1c = item_of[s1];
        const uint32_t grid = std::min<uint32_t>((i1c - i0 + 3) / 4, uint32_t(walk_ctas));
        for (uint32_t pass = 1; pass <= (seg_mode ? 2u : 1u) && i1c > i0; ++pass) {
            seg.pass = pass;
            ask_walk_kernel<<<grid, 128, 0, st>>>(dx, ds, i0, i1c, tab, static_cast<const uint32_t*>(e->d_ask_bits.p), pchk, raw_recs,
                                                  raw_ord, rec_cnt, dev_cap, ctr, ctr + 2 + 2 * g + (pass - 1), seg, raw_item,
                                                  want_bits ? 1u : 0u);
            e->launches += 1;
        }
        CK(cudaGetLastError());
        if (n_chunks > 1) {
            CK(cudaEventRecord(e->wev[g % 4], st));
            CK(cudaStreamWaitEvent(e->stream, e->wev[g % 4], 0));
        }
    }
    CK(cudaEventRecord(e->ev[2], e->stream));
    unsigned long long hc[2] = {0, 0};
    CK(cudaMemcpyAsync(hc, e->d_ask_l2.p, sizeof hc, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    cudaEventElapsedTime(&e->ask_ms[0], e->ev[0], e->ev[1]);
    cudaEventElapsedTime(&e->ask_ms[1], e->ev[1], e->ev[2]);
    *n_recs = hc[0];
    *dx_out = dx;
    *dev_cap_out = dev_cap;
    if (hc[0] > dev_cap) return tmk_fail(e, TMK_E_CAPACITY, "chirp synchroniser: internal record scratch too small");
    if (seg_mode) {
        // which records count: from walker 0 of every capture along the meeting points; then (capture, time) order
        e->launches += tmk_launch_seg_chain(seg.items, seg.st, n_items, first_valid, n_valid, e->stream);
        e->launches += tmk_launch_excl_scan(n_valid, rec_base, n_items, e->stream);
        uint64_t n_exact = 0;
        CK(cudaMemcpyAsync(&n_exact, rec_base + n_items, sizeof n_exact, cudaMemcpyDeviceToHost, e->stream));
        if (hc[0]) {
            ask_seg_order_kernel<<<uint32_t((hc[0] + 255) / 256), 256, 0, e->stream>>>(raw_recs, raw_ord, raw_item, hc[0], rec_base,
                                                                                      first_valid, static_cast<tmk_ask_rec*>(e->d_ask_recs.p));
            e->launches += 1;
        }
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(e->stream));
        *n_recs = n_exact;
        return TMK_OK;
    }
    if (hc[0]) {  // (capture, time) order on the device
        e->launches += tmk_launch_excl_scan(rec_cnt, rec_base, n_streams, e->stream);
        ask_order_kernel<<<uint32_t((hc[0] + 255) / 256), 256, 0, e->stream>>>(raw_recs, raw_ord, hc[0], rec_base,
                                                                              static_cast<tmk_ask_rec*>(e->d_ask_recs.p));
        e->launches += 1;
        CK(cudaGetLastError());
    }
    return TMK_OK;
}

extern "C" {

int tmk_ask_decode_batch(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                         uint32_t n_streams, tmk_ask_rec* recs, uint64_t rec_cap, uint64_t* n_recs) {
    if (!e || (!samples && total_samples) || (!streams && n_streams) || !n_recs || (!recs && rec_cap))
        return tmk_fail(e, TMK_E_INVALID, "tmk_ask_decode_batch: null argument");
    *n_recs = 0;
    if (n_streams == 0) return TMK_OK;
    const float* dx = nullptr;
    uint64_t dev_cap = 0;
    unsigned long long hc[1] = {0};
    {
        uint64_t n = 0;
        const int rc = tmk_ask_sync_device(e, samples, total_samples, streams, n_streams, &dx, &n, &dev_cap);
        hc[0] = n;
        *n_recs = n;
        if (rc != TMK_OK) return rc;
    }
    if (hc[0] > rec_cap) return tmk_fail(e, TMK_E_CAPACITY, "tmk_ask_decode_batch: record buffer too small");
    if (hc[0]) {
        CK(cudaMemcpyAsync(recs, e->d_ask_recs.p, hc[0] * sizeof(tmk_ask_rec), cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
        // (already in reference order — per capture, in time: ask_order_kernel)
    }
    return TMK_OK;
}

// Synthetic chirp + BPSK captures (measurement harness): AWGN of the given SNR against the BPSK carrier power 0.5
//  plus frames laid out like the reference's TX loop, generated on the GPU. samples: device or host.
int tmk_ask_synth_batch(tmk_engine* e, uint64_t seed, uint32_t n_streams, uint64_t n_per_stream, float snr_db, float* samples,
                        uint64_t* n_frames) {
    if (!e || !samples || !n_frames) return tmk_fail(e, TMK_E_INVALID, "tmk_ask_synth_batch: null argument");
    if (n_per_stream % 4 != 0 || n_per_stream >= (1ull << 32))
        return tmk_fail(e, TMK_E_INVALID, "tmk_ask_synth_batch: n_per_stream must be a multiple of 4 below 2^32");
    *n_frames = 0;
    if (n_streams == 0) return TMK_OK;
    CK(cudaSetDevice(e->device));
    int rc = ensure_tables(e);
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
    const float sigma = snr_db >= 100.0f ? 0.0f : std::sqrt(0.5f) * std::pow(10.0f, -snr_db / 20.0f);
    e->launches += tmk_launch_synth_noise(seed ^ 0xA5Cull, total, sigma, n_per_stream, imp, d_x, e->stream);
    CK(e->d_ask_l2.reserve(4 * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(e->d_ask_l2.p, 0, 4 * sizeof(unsigned long long), e->stream));
    e->launches += tmk_launch_ask_synth(static_cast<const float*>(e->d_ask_tab.p), seed, n_streams, n_per_stream, d_x,
                                        static_cast<unsigned long long*>(e->d_ask_l2.p), e->stream);
    CK(cudaGetLastError());
    unsigned long long nf = 0;
    CK(cudaMemcpyAsync(&nf, e->d_ask_l2.p, sizeof nf, cudaMemcpyDeviceToHost, e->stream));
    if (!dev_out) CK(cudaMemcpyAsync(samples, d_x, total * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    *n_frames = nf;
    return TMK_OK;
}

int tmk_ask_last_stage_ms(tmk_engine* e, float ms[2]) {
    if (!e || !ms) return TMK_E_INVALID;
    ms[0] = e->ask_ms[0];
    ms[1] = e->ask_ms[1];
    return TMK_OK;
}

}  // extern "C"
