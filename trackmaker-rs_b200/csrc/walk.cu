// walk.cu — per-stream walk: turns the scan kernel's candidate bitmap into the exact sequence of
// verdicts the reference decoder produces (src/phy/decoder.rs:68-356), one warp per stream.
//
// The walk emulates the PhyDecoder state machine over *sparse* events only:
//   Searching  (decoder.rs:123-238): the first window with corr >= threshold, where corr uses the
//              reference's AVX-order dot product (:381-417) and its ROLLING window energy
//              (:224-233), re-seeded at every search entry. The rolling value E_ref is not
//              computed; instead  E_ref in [E_c - delta, E_c + delta]  with E_c an accurately
//              summed window energy and delta a rigorous bound on the rolling drift since the
//              seed (see DESIGN.md "exactness"). Because corr is monotone in E, evaluating the
//              reference formula at both ends decides almost every window; the rare undecided
//              one is replayed sequentially from its seed (TMK_STAT_EXACT_REPLAYS).
//   refinement (:157-221), Decoding (:242-356), line decode (line_coding.rs:93-123,243-293),
//              header parse / CRC (frame.rs:109-183, crc.rs:7-22): executed literally, in the
//              reference's f32 operation order, warp-parallel over offsets / levels / bytes.
// Chunked feeding (cfg.chunk) is emulated through `avail`, the number of samples the reference
// would have buffered when it examines a window.
#include "tmk_internal.cuh"

namespace {

#ifndef TMK_CONTRAST
#define TMK_CONTRAST 512  // max / min window energy of the block that holds q beyond which its exact energies are computed
#endif
constexpr int kWarpsPerCta = 4;
constexpr int kMaxFrameBytesCap = 1024;                 // cfg.max_frame_bytes upper limit
constexpr int kByteBuf = TMK_PHY_HEADER_BYTES + kMaxFrameBytesCap + 9;
constexpr int kFastMaxP = 127;                          // longest preamble of the staged-block paths (shipped shapes: 60, 96)
constexpr int kQBuf = TMK_BLK + kFastMaxP + 1;          // prefix sums of one block for P <= kFastMaxP
constexpr int kXsMaxP = 96;                             // longest preamble whose block samples are staged (shipped shapes: 60, 96)
constexpr int kXs = TMK_BLK + kXsMaxP;
constexpr float kU = 5.9604645e-8f;                     // 2^-24, unit roundoff of f32
constexpr unsigned kFull = 0xffffffffu;
// the scan kernels' window energies (fp32 FMA chains of at most P + 2 x 64 roundings from a strip / row start) are
// within kScanErr x (largest window energy of the neighbourhood) of the exact sums
constexpr float kScanErr = 256.f * kU;
// A window outside the candidate bitmap has corr_approx < thr * TMK_CAND_RATIO; the reference can
// only detect it if its rolling energy sits below (TMK_CAND_RATIO)^2 = 0.892 of the true window
// energy. kTight is the share of the window energy the drift bound may use up before such a
// window has to be examined individually (0.09 < 1 - 0.892, the rest is margin for the scan
// kernel's own rounding).
constexpr float kTight = 0.09f;

__device__ uint8_t g_crc_pow[kMaxFrameBytesCap + 8];    // x^(8k) mod 0x107, k = 0..

struct alignas(16) WarpScratch {
    alignas(16) uint8_t bytes[kByteBuf];  // decoded frame bytes; staging of the replay during a search
    double Q[kQBuf + 1];
    int8_t pre[TMK_MAX_P];  // preamble samples (+1/-1 levels expanded by spl; x * +-1.0f is exact)
    alignas(16) float xs[kXs];  // the samples of the block whose prefix sums are in Q (block_prefix), for P <= kXsMaxP
};
static_assert(sizeof(WarpScratch) * kWarpsPerCta + 64 * 16 * kWarpsPerCta + 256 <= 48 * 1024, "static shared memory of the walk / cluster kernels");

// One scratch per warp, at namespace scope so that every access compiles to LDS / STS (through a pointer
// handed from function to function the accesses are generic loads: slower, and on the long scoreboard).
__shared__ WarpScratch g_scratch[kWarpsPerCta];

struct ScratchRef {
    uint32_t w;
    __device__ __forceinline__ WarpScratch* operator->() const { return &g_scratch[w]; }
};

struct Ctx {
    const float* x;       // stream samples
    uint64_t N;           // samples in the stream
    uint64_t nwin;        // windows (N - P + 1, or 0)
    uint64_t C;           // chunk (0 = one shot)
    const TmkPattern* pat;
    const float4* blk_info;
    const uint32_t* cand_bits;
    const uint2* blk_res;            // cluster pass results of this stream's blocks (nullptr: none)
    const TmkCandEntry* entries;
    const tmk_frame_rec* spec;
    uint64_t nblk;
    float thr;
    int32_t local_addr;
    uint32_t max_frame_bytes;
    uint32_t P;
    int lane;
    ScratchRef sc;
    // exact replay cache: rolling energy `rE` of window `rpos` for a chain seeded at `rseed`
    uint64_t rseed, rpos;
    float rE;
    bool rvalid;
    // anchor: |E_ref(a_w) - S(a_w)| <= a_err for the chain seeded at a_seed (a_w == a_seed, a_err == 0
    // right after a search entry; a measured value after an exact re-anchoring)
    uint64_t a_seed, a_w;
    uint64_t n_par_steps;  // steps rolled by the lane-parallel replay, and segments it had to roll literally
    uint32_t n_par_doubt;
    float a_err;
    // sum form of the drift bound, valid while the anchor is the seed of a one-shot search (see fast_block):
    // bound on the sum of the chain's window energies before the block handed to process_block, and on
    // the energy of the seed window
    float racc_in, efirst;
    bool sum_ok;
    // stats
    uint32_t n_replays, n_cands, n_det;
    // diagnostics: [0] process_block calls, [1] re-anchors, [2] replay steps, [3] blk_info groups
    unsigned long long dbg[4];
};

// samples the reference has buffered when it examines window w, given the current `avail`
__device__ __noinline__ uint64_t avail_for_slow(const uint64_t P, const uint64_t C, const uint64_t N, uint64_t w) {
    uint64_t t = ((w + P + C - 1) / C) * C;  // C > 0 here (one shot has avail == N)
    return t < N ? t : N;
}

__device__ __forceinline__ uint64_t avail_for(const Ctx& c, uint64_t w, uint64_t avail) {
    if (w + c.P <= avail) return avail;
    return avail_for_slow(c.P, c.C, c.N, w);
}

// position at which the search that examines window w was (re)entered = where its rolling
// energy was seeded (decoder.rs:133-136)
__device__ __noinline__ uint64_t seed_for_slow(const uint64_t P, const uint64_t C, const uint64_t N, uint64_t w, uint64_t q) {
    uint64_t aw = avail_for_slow(P, C, N, w);
    uint64_t prev = ((aw - 1) / C) * C;  // avail of the previous process_samples call
    if (prev + 1 < P) return q;
    uint64_t s = prev + 1 - P;
    return s > q ? s : q;
}

__device__ __forceinline__ uint64_t seed_for(const Ctx& c, uint64_t w, uint64_t q, uint64_t avail) {
    if (w + c.P <= avail) return q;
    return seed_for_slow(c.P, c.C, c.N, w, q);
}

// w * p for a preamble sample p = +-1.0f: exactly w with the sign flipped for p = -1 (what the f32 multiply returns,
// signed zeros and NaN payloads included), without the int -> float conversion and the multiply
__device__ __forceinline__ float mul_pm1(float w, int8_t p) {
    return __uint_as_float(__float_as_uint(w) ^ (uint32_t(int32_t(p)) & 0x80000000u));
}

// compute_dot_product_avx (decoder.rs:381-417): 8 lanes of mul-then-add, lane l + lane l+4,
// the 4 partials summed left to right from 0, then the len % 8 tail. Runs per thread.
__device__ __noinline__ float dot_avx_order(const Ctx& c, const float* __restrict__ w) {
    const uint32_t P = c.P;
    const int8_t* __restrict__ pre = c.sc->pre;
    float a[8];
#pragma unroll
    for (int l = 0; l < 8; ++l) a[l] = 0.0f;
    uint32_t i = 0;
    for (; i + 8 <= P; i += 8) {
#pragma unroll
        for (int l = 0; l < 8; ++l) a[l] = ex_add(a[l], mul_pm1(w[i + l], pre[i + l]));
    }
    float s = 0.0f;
#pragma unroll
    for (int l = 0; l < 4; ++l) s = ex_add(s, ex_add(a[l], a[l + 4]));
    for (; i < P; ++i) s = ex_add(s, mul_pm1(w[i], pre[i]));
    return s;
}

// The same dot product for up to 4 windows at once, out of the block's staged samples (c.sc->xs, block_prefix):
// the 8 lanes of group g = lane / 8 ARE the 8 AVX lanes of window wl[g], so the dependent chain is P / 8 adds
// instead of P. wl: window offset inside the block (warp-uniform per group; groups beyond `n` idle).
// Returns the dot product of the group's window on every lane of the group.
__device__ __forceinline__ float dot_avx_coop(const Ctx& c, uint32_t wl_g, bool active) {
    const uint32_t P = c.P;
    const int l = c.lane & 7;
    const int8_t* __restrict__ pre = c.sc->pre;
    const float* __restrict__ w = c.sc->xs + wl_g;
    float a = 0.0f;
    uint32_t i = 0;
    if (active)
        for (; i + 8 <= P; i += 8) a = ex_add(a, mul_pm1(w[i + l], pre[i + l]));
    i = P & ~7u;
    // lane l + lane l + 4 (l < 4), then the four partials left to right from 0
    const float hi = __shfl_down_sync(kFull, a, 4);
    const float pr = ex_add(a, hi);
    const int base = c.lane & ~7;
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < 4; ++k) s = ex_add(s, __shfl_sync(kFull, pr, base + k));
    if (active)
        for (; i < P; ++i) s = ex_add(s, mul_pm1(w[i], pre[i]));
    return s;
}

// iter().map(|x| x * x).sum::<f32>()  — sequential left to right
__device__ float seq_energy(const float* __restrict__ w, uint32_t n) {
    float e = 0.0f;
    for (uint32_t i = 0; i < n; ++i) e = ex_add(e, ex_mul(w[i], w[i]));
    return e;
}

// The literal rolling chain from position `pos` (value e) to window `w` (decoder.rs:224-233). Sequential by
// construction: the squares are formed lane-parallel, the dependent (sub, add, clamp) chain runs redundantly
// on every lane.
__device__ __noinline__ float roll_literal(Ctx& c, float e, uint64_t pos, uint64_t w) {
    const float* __restrict__ x = c.x;
    const uint32_t P = c.P;
    c.dbg[2] += w - pos;
    // Full groups of kG x 32 steps. The loads of the next group are in flight while the current one
    // is consumed (the chain is a dependent add/add per step; without the prefetch every 32 steps
    // paid one exposed HBM round trip). The squares of a group are staged in shared memory (the
    // byte buffer of the frame decoder is idle during a search) and read back as broadcast float4s.
    // A group of 128 steps first runs WITHOUT the clamp `if e < 0 { e = 0 }` (decoder.rs:231-233),
    // tracking the minimum off the critical path: if no intermediate value was negative the clamp
    // never fired and the values are the reference's; otherwise the group is redone literally.
    // NaN/inf need no special case: they propagate identically with and without the clamp.
    constexpr int kG = 4;
    constexpr int kQuads = 8 * kG;  // float4s of a group per operand
    static_assert(kByteBuf >= 2 * 32 * kG * 4, "staging buffer of the replay");
    if (pos + 32 * kG <= w) {
        float* sa = reinterpret_cast<float*>(c.sc->bytes);  // a[128] | b[128]
        float* sb = sa + 32 * kG;
        // global loads run TWO groups ahead of the chain (one group is ~1.5 K cycles of chain, less than a
        // loaded HBM round trip), shared-memory loads two float4 pairs ahead: the warp issues in order, so
        // anything the chain has to wait for stalls all of it
        float lv[kG], en[kG], lv2[kG], en2[kG];
#pragma unroll
        for (int g = 0; g < kG; ++g) {
            lv[g] = x[pos + g * 32 + c.lane];
            en[g] = x[pos + P + g * 32 + c.lane];
            lv2[g] = en2[g] = 0.f;
        }
        if (pos + 2 * 32 * kG <= w) {
#pragma unroll
            for (int g = 0; g < kG; ++g) {
                lv2[g] = x[pos + 32 * kG + g * 32 + c.lane];
                en2[g] = x[pos + 32 * kG + P + g * 32 + c.lane];
            }
        }
        for (;;) {
            const bool more = pos + 2 * 32 * kG <= w;
            __syncwarp();
#pragma unroll
            for (int g = 0; g < kG; ++g) {
                sa[g * 32 + c.lane] = ex_mul(lv[g], lv[g]);
                sb[g * 32 + c.lane] = ex_mul(en[g], en[g]);
            }
            __syncwarp();
#pragma unroll
            for (int g = 0; g < kG; ++g) {
                lv[g] = lv2[g];
                en[g] = en2[g];
            }
            if (pos + 3 * 32 * kG <= w) {
#pragma unroll
                for (int g = 0; g < kG; ++g) {
                    lv2[g] = x[pos + 2 * 32 * kG + g * 32 + c.lane];
                    en2[g] = x[pos + 2 * 32 * kG + P + g * 32 + c.lane];
                }
            }
            const float4* a4 = reinterpret_cast<const float4*>(sa);
            const float4* b4 = reinterpret_cast<const float4*>(sb);
            const float e0 = e;
            float mn = e;
            float4 A[3], B[3];
            A[0] = a4[0]; B[0] = b4[0];
            A[1] = a4[1]; B[1] = b4[1];
#pragma unroll
            for (int j = 0; j < kQuads; ++j) {
                if (j + 2 < kQuads) {
                    A[(j + 2) % 3] = a4[j + 2];
                    B[(j + 2) % 3] = b4[j + 2];
                }
                const float4 a = A[j % 3], b = B[j % 3];
                e = ex_add(ex_sub(e, a.x), b.x); mn = fminf(mn, e);
                e = ex_add(ex_sub(e, a.y), b.y); mn = fminf(mn, e);
                e = ex_add(ex_sub(e, a.z), b.z); mn = fminf(mn, e);
                e = ex_add(ex_sub(e, a.w), b.w); mn = fminf(mn, e);
            }
            if (!(mn >= 0.0f)) {  // a clamp fired (or e0 was NaN): literal form
                e = e0;
#pragma unroll 1
                for (int j = 0; j < 32 * kG; ++j) {
                    e = ex_add(ex_sub(e, sa[j]), sb[j]);
                    e = e < 0.0f ? 0.0f : e;
                }
            }
            pos += 32 * kG;
            if (!more) break;
        }
        __syncwarp();
    }
    while (pos < w) {
        const uint64_t left = w - pos;
        const uint32_t n = left < 32 ? uint32_t(left) : 32u;
        float lv = 0.f, en = 0.f;
        if (uint32_t(c.lane) < n) {
            lv = x[pos + c.lane];
            en = x[pos + P + c.lane];
        }
        lv = ex_mul(lv, lv);
        en = ex_mul(en, en);
        for (uint32_t j = 0; j < n; ++j) {
            e = ex_add(ex_sub(e, __shfl_sync(kFull, lv, j)), __shfl_sync(kFull, en, j));
            e = e < 0.0f ? 0.0f : e;
        }
        pos += n;
    }
    return e;
}

// ---------------------------------------------------------------- lane-parallel exact replay
// The chain  e' = clamp(fl(fl(e - a) + b))  is serial, but it is TRANSLATION INVARIANT inside a binade: if every
// intermediate t of a chain started at g' and its translate t + s lie in the same binade [2^m, 2^(m+1)), and s is
// an even multiple of the coarsest rounding grid that occurs, then every rounding (ties included: an even
// multiple of the grid keeps the parity of both neighbours) and therefore the whole chain commutes with the
// translation:  chain(g' + s) = chain(g') + s,  exactly. So the 32 lanes roll 32 consecutive segments at once,
// each from a GUESS g (the accurately summed window energy at the segment start) — in fact from the four bases
// g, g+u, g+2u, g+3u (u = ulp(g)), so that whatever the true start e is, one base differs from it by a multiple
// of 4u, which is an even multiple of the coarsest grid as long as the segment stays at most one binade above g.
// The step map is monotone, so the four chains stay ordered; the smallest distance of chain 0 to the binade
// boundary below and of chain 3 to the boundary above bounds the margin of all four. The segments are then
// stitched in order: true start - base = s, accepted if |s| is inside the segment's margin (in units of its
// smallest ulp, one ulp held back for the rounded vs real intermediate), negative / zero / subnormal / NaN /
// inf values, a clamp, or chains 0 and 3 in different binades make the segment "in doubt": it is rolled
// literally (roll_literal) from the exact start value that is known by then.
#ifndef TMK_PAR_DEPTH
#define TMK_PAR_DEPTH 1  // iterations (of 8 steps) the sample loads of the lane-parallel replay run ahead of the chain
                         // (2 measured on the final kernel: walk 2.92 -> 2.95 ms at C2, 1.43 -> 1.41 ms on 16 384 x 120 000: no gain)
#endif
#ifndef TMK_PAR_MIN
#define TMK_PAR_MIN 128
#endif
constexpr uint32_t kParMinSeg = TMK_PAR_MIN;  // steps per lane below which a parallel round is not worth it
#ifndef TMK_PAR_SEG
#define TMK_PAR_SEG 512
#endif
constexpr uint32_t kParMaxSeg = TMK_PAR_SEG;  // steps per lane and round: the chance that a segment is in doubt and the cost of
                                              // rolling it literally both grow with it (2048: 42 % of the segments at C2)
                                              // (final kernel, walk at C2, ms: 256 -> 2.956, 384 -> 2.924, 512 -> 2.916, 768 -> 3.002,
                                              //  1024 -> 3.064; TMK_PAR_MIN 64 -> 2.912, 128 -> 2.916, 256 -> 2.979)

__device__ __noinline__ float roll_parallel(Ctx& c, float e, uint64_t* pos_io, uint64_t w) {
    const float* __restrict__ x = c.x;
    const uint32_t P = c.P;
    const int lane = c.lane;
    constexpr uint32_t kM = 0x7fffffu;
    uint64_t pos = *pos_io;
    while (w - pos >= 32ull * kParMinSeg) {
        const uint64_t l64 = (w - pos) / 32;
        const uint32_t L = uint32_t(l64 < kParMaxSeg ? l64 : kParMaxSeg) & ~7u;
        const float* __restrict__ xp = x + pos + uint64_t(lane) * L;  // this lane's segment
        // guess: accurately summed energy of the window at the segment start
        double S = 0.0;
        for (uint32_t i = 0; i < P; i += 4) {
            const float4 v = *reinterpret_cast<const float4*>(xp + i);
            S += double(ex_mul(v.x, v.x));
            S += double(ex_mul(v.y, v.y));
            S += double(ex_mul(v.z, v.z));
            S += double(ex_mul(v.w, v.w));
        }
        // The chain's drift against the accurately summed energy is a slow random walk (hundreds of ulps after 100 k
        // steps, a few tens more over one round), so every guess is shifted by the drift measured at the start of
        // the round: the translations that remain are a few tens of ulps and far fewer segments end up in doubt.
        float g = __double2float_rn(S);
        {
            const float d0 = ex_sub(e, __shfl_sync(kFull, g, 0));
            if (fabsf(d0) <= 0.01f * g) g = ex_add(g, d0);  // (NaN / wild values: keep the plain guess)
        }
        const uint32_t gbits = __float_as_uint(g);
        const uint32_t gexp = gbits >> 23;  // (sign bit set -> huge -> rejected)
        bool ok = gexp >= 32u && gexp <= 250u && (gbits & kM) <= 0x7ffff0u;
        const float u = __uint_as_float(((ok ? gexp : 64u) - 23u) << 23);  // ulp(g)
        float c0 = g, c1 = ex_add(g, u), c2 = ex_add(g, ex_add(u, u)), c3 = ex_add(g, ex_mul(3.0f, u));
        uint32_t below = kM, top = 0u, xr = 0u, hi = 0u;
        int32_t lo = 0x7fffffff;
        // samples as float4s (pos and P are multiples of 4 here, the stream base is 16-byte aligned): every load
        // instruction touches 32 different lines, so the instruction count is what the L1 pays for.
        // (Tried in round 2: the samples through a per-lane cp.async ring in shared memory, four iterations ahead of
        //  the chain — the 43 KB of dynamic shared memory per CTA come out of the L1 the rest of the walk lives on:
        //  walk 2.9 -> 4.5 ms at C2. An L1 prefetch of the lines two ahead changed nothing.)
        const float4* __restrict__ xa = reinterpret_cast<const float4*>(xp);
        const float4* __restrict__ xb = reinterpret_cast<const float4*>(xp + P);
        float4 a0 = xa[0], a1 = xa[1], b0 = xb[0], b1 = xb[1];
#if TMK_PAR_DEPTH == 2
        float4 pa0 = a0, pa1 = a1, pb0 = b0, pb1 = b1;  // the samples of the next iteration: loads run two iterations ahead
        if (L > 8) {
            pa0 = xa[2];
            pa1 = xa[3];
            pb0 = xb[2];
            pb1 = xb[3];
        }
#endif
        for (uint32_t k = 0; k < L; k += 8) {
#if TMK_PAR_DEPTH == 2
            float4 an0 = pa0, an1 = pa1, bn0 = pb0, bn1 = pb1;
            if (k + 16 < L) {
                pa0 = xa[k / 4 + 4];
                pa1 = xa[k / 4 + 5];
                pb0 = xb[k / 4 + 4];
                pb1 = xb[k / 4 + 5];
            }
#else
            float4 an0 = a0, an1 = a1, bn0 = b0, bn1 = b1;
            if (k + 8 < L) {
                an0 = xa[k / 4 + 2];
                an1 = xa[k / 4 + 3];
                bn0 = xb[k / 4 + 2];
                bn1 = xb[k / 4 + 3];
            }
#endif
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const float aa = ex_mul(a[t], a[t]), bb = ex_mul(b[t], b[t]);
                const float t0 = ex_sub(c0, aa), t1 = ex_sub(c1, aa), t2 = ex_sub(c2, aa), t3 = ex_sub(c3, aa);
                c0 = ex_add(t0, bb);
                c1 = ex_add(t1, bb);
                c2 = ex_add(t2, bb);
                c3 = ex_add(t3, bb);
                const uint32_t bt0 = __float_as_uint(t0), bt3 = __float_as_uint(t3);
                const uint32_t bc0 = __float_as_uint(c0), bc3 = __float_as_uint(c3);
                below = min(below, min(bt0 & kM, bc0 & kM));
                top = max(top, max(bt3 & kM, bc3 & kM));
                xr |= (bt0 ^ bt3) | (bc0 ^ bc3);
                lo = min(lo, min(int32_t(bt0), int32_t(bc0)));
                hi = max(hi, max(bt3, bc3));
            }
            a0 = an0;
            a1 = an1;
            b0 = bn0;
            b1 = bn1;
        }
        // chains 0 and 3 in the same binade (and of the same sign) at every operation; everything normal and
        // positive (no clamp, no gate); at most one binade above the guess
        ok = ok && (xr >> 23) == 0u && lo >= int32_t(32u << 23) && (hi >> 23) <= gexp + 1u && (hi >> 23) <= 250u;
        float margin = 0.f;
        if (ok) {
            const uint32_t mi = min(below, 0x800000u - top);  // in ulps of the value it was measured on (>= smallest ulp)
            const uint32_t emin = uint32_t(lo) >> 23;
            if (mi >= 2u) margin = ex_mul(float(mi - 1u), __uint_as_float((emin - 23u) << 23));
        }
        const float inv_u = __uint_as_float((277u - (ok ? gexp : 64u)) << 23);  // 1 / ulp(g), exact
        // ---- stitch the segments in order (warp-uniform)
        float ecur = e;
        for (int j = 0; j < 32; ++j) {
            const bool okj = __shfl_sync(kFull, int(ok), j) != 0;
            const float gj = __shfl_sync(kFull, g, j), uj = __shfl_sync(kFull, u, j), iuj = __shfl_sync(kFull, inv_u, j);
            const float mj = __shfl_sync(kFull, margin, j);
            const float r0 = __shfl_sync(kFull, c0, j), r1 = __shfl_sync(kFull, c1, j);
            const float r2 = __shfl_sync(kFull, c2, j), r3 = __shfl_sync(kFull, c3, j);
            bool use = okj && (__float_as_uint(ecur) >> 23) == (__float_as_uint(gj) >> 23);  // same binade, positive
            float enext = 0.f;
            if (use) {
                const float d = ex_sub(ecur, gj);   // exact: same binade
                const float qf = ex_mul(d, iuj);    // exact: a power of two
                use = fabsf(qf) < 4.0e6f;
                if (use) {
                    const int iq = __float2int_rn(qf);  // qf is an integer
                    const int i = iq & 3;
                    const float sft = ex_mul(float(iq - i), uj);  // multiple of 4 ulp(g), exact
                    use = fabsf(sft) <= mj;
                    if (use) enext = ex_add(i == 0 ? r0 : i == 1 ? r1 : i == 2 ? r2 : r3, sft);  // exact by the lemma
                }
            }
            if (!use) {
                enext = roll_literal(c, ecur, pos + uint64_t(j) * L, pos + uint64_t(j + 1) * L);
                c.n_par_doubt++;
            }
            ecur = enext;
        }
        c.n_par_steps += 32ull * L;
        e = ecur;
        pos += 32ull * L;
    }
    *pos_io = pos;
    return e;
}

// Exact rolling window energy of window `w` for a search seeded at `seed` (decoder.rs:133-136, 224-233).
__device__ __noinline__ float exact_energy(Ctx& c, uint64_t seed, uint64_t w) {
    float e;
    uint64_t pos;
    if (c.rvalid && c.rseed == seed && c.rpos <= w) {
        e = c.rE;
        pos = c.rpos;
    } else {
        e = seq_energy(c.x + seed, c.P);
        pos = seed;
    }
    if (w - pos >= 32ull * kParMinSeg + 4 && c.P % 4 == 0) {
        const uint64_t al = (pos + 3) & ~3ull;  // the parallel replay loads float4s
        if (al != pos) e = roll_literal(c, e, pos, al);
        pos = al;
        e = roll_parallel(c, e, &pos, w);
    }
    e = roll_literal(c, e, pos, w);
    c.rvalid = true;
    c.rseed = seed;
    c.rpos = w;
    c.rE = e;
    c.n_replays++;
    return e;
}

// Bound on |E_ref(w) - S(w)|, S = exact sum of the rounded squares of window w, for a chain that
// has run k rolling steps since a point where the error was at most e0, with every window energy
// on the way at most smax (DESIGN.md "exactness"): each step adds at most 2.1 u S and multiplies
// the error by at most (1 + 2.1 u). The 1.5 P u smax term covers the sequential seed sum.
__device__ __forceinline__ float drift_bound(uint32_t P, uint64_t k, float e0, float smax) {
    const float kf = float(k);
    const float growth = kf < 65536.f ? 1.01f : __expf(2.2f * kU * kf) * 1.01f;
    return (e0 + kU * (1.5f * float(P) + 2.2f * kf) * smax) * growth;
}

enum { DEC_NO = 0, DEC_YES = 1, DEC_AMBIG = 2 };

// decide corr_ref >= thr (thr > 0) for a window with exact dot D and E_ref in [Ec-delta, Ec+delta].
// corr = D / (sqrt(E) * pe) is monotone in E under round-to-nearest, and the energy gate
// (E < 1e-6 -> corr = 0, decoder.rs:142-144) can only turn a detection into a non-detection.
__device__ __forceinline__ int decide_interval(const Ctx& c, float D, float Ec, float delta) {
    if (!(D > 0.f)) return DEC_NO;  // corr <= 0, gated 0, or NaN: never >= thr
    if (!(delta >= 0.f) || !(delta <= 3.0e38f) || !(Ec >= 0.f) || !(Ec <= 3.0e38f) || !(D <= 3.0e38f))
        return DEC_AMBIG;  // NaN / inf somewhere on the chain: let the literal replay decide
    float lo = __fsub_rd(Ec, delta), hi = __fadd_ru(Ec, delta);
    if (!(hi <= 3.0e38f)) return DEC_AMBIG;
    if (hi < 1e-6f) return DEC_NO;  // gated whatever E_ref is
    const float lo_g = lo < 1e-6f ? 1e-6f : lo;
    if (ex_corr(D, lo_g, c.pat->pe) < c.thr) return DEC_NO;  // even the smallest ungated energy fails
    if (lo >= 1e-6f && ex_corr(D, hi, c.pat->pe) >= c.thr) return DEC_YES;
    return DEC_AMBIG;
}

__device__ __forceinline__ double warp_incl_scan_add(double v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        double t = __shfl_up_sync(kFull, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

__device__ __forceinline__ float warp_incl_scan_maxnan(float v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        float t = __shfl_up_sync(kFull, v, o);
        if (lane >= o) v = max_nan(v, t);
    }
    return v;
}

// accurately summed energy of window w (double accumulation of the rounded squares), warp-parallel
__device__ float window_energy(const Ctx& c, uint64_t w) {
    double s = 0.0;
    for (uint32_t k = c.lane; k < c.P; k += 32) {
        const float v = c.x[w + k];
        s += double(ex_mul(v, v));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(kFull, s, o);
    return __double2float_rn(s);
}

struct Detection {
    uint64_t c;      // detecting window (preamble start)
    uint64_t seed;   // where the search that found it was entered (search area start)
    uint64_t avail;  // samples buffered at that time
    float corr;      // reference correlation value of the detecting window
    uint32_t spec;   // speculative verdict record of the cluster pass for this window, or TMK_ENT_INVALID
    // what the walk needs of that record to go on (from the candidate entry; status == TMK_ENT_NOSPEC: read the record)
    uint32_t status, consumed, len;
    uint64_t fs;
};

// Exact evaluation of window w (literal reference arithmetic). All lanes, uniform arguments.
__device__ bool exact_window(Ctx& c, uint64_t w, uint64_t seed, float D, float* corr_out) {
    float e = exact_energy(c, seed, w);
    float corr = ex_corr(D, e, c.pat->pe);
    *corr_out = corr;
    return corr >= c.thr;
}

// Prefix sums of the rounded squares of block `bb` in double: E_c(w) = Q[w+P] - Q[w]  (P <= uint32_t(kFastMaxP)).
// Every lane owns K consecutive samples, fetched as float4s that are all in flight together (one HBM
// round trip for the block), summed serially per lane, then one warp scan of the 32 lane totals.
__device__ __noinline__ void block_prefix(Ctx& c, uint64_t bb) {
    const int lane = c.lane;
    const uint32_t P = c.P;
    const uint64_t wb0 = bb * TMK_BLK;
    const float* __restrict__ x = c.x;
    {

        // prefix sums of the rounded squares in double: E_c(w) = Q[w+P] - Q[w].
        // Every lane owns K consecutive samples, fetched as float4s that are all in flight together
        // (one HBM round trip for the block), summed serially per lane, then one warp scan of the
        // 32 lane totals.
        uint64_t ns64 = c.N - wb0;
        const uint32_t ns = ns64 < uint64_t(TMK_BLK + P - 1) ? uint32_t(ns64) : TMK_BLK + P - 1;
        constexpr int kMaxK4 = (TMK_BLK + kFastMaxP + 127) / 128;  // float4s per lane
        const uint32_t K4 = (ns + 127) / 128;
        const uint32_t i0 = uint32_t(lane) * K4 * 4;
        const float* __restrict__ xb = x + wb0 + i0;
        float4 v[kMaxK4];
#pragma unroll
        for (int k = 0; k < kMaxK4; ++k) {
            v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (uint32_t(k) < K4) {
                const uint32_t i = i0 + 4 * k;
                if (i + 4 <= ns) {
                    v[k] = *reinterpret_cast<const float4*>(xb + 4 * k);
                } else {
                    if (i + 0 < ns) v[k].x = xb[4 * k + 0];
                    if (i + 1 < ns) v[k].y = xb[4 * k + 1];
                    if (i + 2 < ns) v[k].z = xb[4 * k + 2];
                }
            }
        }
        double tot = 0.0;
#pragma unroll
        for (int k = 0; k < kMaxK4; ++k) {
            tot += double(ex_mul(v[k].x, v[k].x));
            tot += double(ex_mul(v[k].y, v[k].y));
            tot += double(ex_mul(v[k].z, v[k].z));
            tot += double(ex_mul(v[k].w, v[k].w));
        }
        double run = warp_incl_scan_add(tot, lane) - tot;  // exclusive
        double* Q = c.sc->Q;
        if (lane == 0) Q[0] = 0.0;
        {   // the samples themselves (candidate dot products read them from here)
            float* xs = c.sc->xs;
#pragma unroll
            for (int k = 0; k < kMaxK4; ++k)
                if (uint32_t(k) < K4 && i0 + 4 * k + 4 <= uint32_t(kXs)) *reinterpret_cast<float4*>(xs + i0 + 4 * k) = v[k];
        }
#pragma unroll
        for (int k = 0; k < kMaxK4; ++k) {
            if (uint32_t(k) < K4) {
                const uint32_t i = i0 + 4 * k;
                run += double(ex_mul(v[k].x, v[k].x));
                if (i + 0 < ns) Q[i + 1] = run;
                run += double(ex_mul(v[k].y, v[k].y));
                if (i + 1 < ns) Q[i + 2] = run;
                run += double(ex_mul(v[k].z, v[k].z));
                if (i + 2 < ns) Q[i + 3] = run;
                run += double(ex_mul(v[k].w, v[k].w));
                if (i + 3 < ns) Q[i + 4] = run;
            }
        }
        __syncwarp();
    }
}

// ---- results of the cluster pass (it has finished when the walk of its chunk starts)
// blk_res of a block with candidates: x = (first entry << 6) | count or TMK_ENT_INVALID, y = gain. Blocks without
// candidates (blk_info.w == 0) have no result and are never looked up.
__device__ __forceinline__ TmkCandEntry ld_entry(const TmkCandEntry* p) {
    static_assert(sizeof(TmkCandEntry) == 32, "entry size");
    const uint4 v = __ldcg(reinterpret_cast<const uint4*>(p)), u = __ldcg(reinterpret_cast<const uint4*>(p) + 1);
    TmkCandEntry e;
    e.wl = uint16_t(v.x & 0xffffu);
    e.status = uint16_t(v.x >> 16);
    e.D = __uint_as_float(v.y);
    e.Ec = __uint_as_float(v.z);
    e.spec = v.w;
    e.consumed = u.x;
    e.fs_off = int32_t(u.y);
    e.len = u.z;
    e.pad_ = 0;
    return e;
}
__device__ __forceinline__ tmk_frame_rec ld_spec(const tmk_frame_rec* p) {
    static_assert(sizeof(tmk_frame_rec) == 48, "record size");
    union { tmk_frame_rec r; uint4 v[3]; } u;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    u.v[0] = __ldcg(q);
    u.v[1] = __ldcg(q + 1);
    u.v[2] = __ldcg(q + 2);
    return u.r;
}

// speculative verdict record the cluster pass prepared for window wl of block bb (or TMK_ENT_INVALID)
__device__ __noinline__ uint32_t lookup_spec(const Ctx& c, uint64_t bb, uint32_t wl) {
    if (c.blk_res == nullptr || c.P > uint32_t(kFastMaxP)) return TMK_ENT_INVALID;
    if (c.blk_info[bb].w == 0.f) return TMK_ENT_INVALID;  // not on the cluster pass's list: no result
    const uint32_t ptr = __ldcg(c.blk_res + bb).x;
    if (ptr == TMK_ENT_INVALID) return TMK_ENT_INVALID;
    const uint32_t n = ptr & 63u;
    for (uint32_t r0 = 0; r0 < n; r0 += 32) {
        uint32_t spec = TMK_ENT_INVALID;
        bool match = false;
        if (r0 + uint32_t(c.lane) < n) {
            const TmkCandEntry e = ld_entry(c.entries + (ptr >> 6) + r0 + c.lane);
            match = e.wl == wl;
            spec = e.spec;
        }
        const uint32_t m = __ballot_sync(kFull, match);
        if (m) return __shfl_sync(kFull, spec, __ffs(m) - 1);
    }
    return TMK_ENT_INVALID;
}

// Examine the windows of block `bb` that are >= q, in order. `prevmax` bounds the window energies
// of the chain between the anchor and this block; *pm_out receives the bound including this block.
// Returns true with *det filled on a detection.
__device__ __noinline__ bool process_block(Ctx& c, uint64_t bb, uint64_t q, uint64_t avail, float prevmax, Detection* det,
                              float* pm_out) {
    const int lane = c.lane;
    const uint32_t P = c.P;
    const uint64_t wb0 = bb * TMK_BLK;
    c.dbg[0]++;
    const float* __restrict__ x = c.x;
    const float silent = 0.5f * c.thr * c.thr * 1e-6f;
    // neighbourhood maximum for the scan kernel's own approximation error
    float nb = c.blk_info[bb].y;
    if (bb > 0) nb = max_nan(nb, c.blk_info[bb - 1].y);
    if (bb + 1 < c.nblk) nb = max_nan(nb, c.blk_info[bb + 1].y);
    const float dscan = kScanErr * nb;

    // the block's candidate words: lane r holds word r (one load round instead of 15)
    const uint32_t myword = lane < TMK_BLK_WORDS ? c.cand_bits[bb * TMK_BLK_WORDS + lane] : 0u;
    const bool fast = P <= uint32_t(kFastMaxP);
    if (fast) block_prefix(c, bb);
    float pmcarry = prevmax * 1.001f;  // block maxima come from the approximate scan
    float sumcarry = c.racc_in;
#pragma unroll 1
    for (uint32_t r = 0; r < TMK_BLK_WORDS; ++r) {
        const uint32_t wl = r * 32 + lane;
        const uint64_t w = wb0 + wl;
        const bool valid = w >= q && w < c.nwin;
        if (!__any_sync(kFull, valid)) continue;
        float Ec = 0.f;
        if (valid) {
            if (fast) {
                Ec = __double2float_rn(c.sc->Q[wl + P] - c.sc->Q[wl]);
            } else {
                double s = 0.0;
                for (uint32_t k = 0; k < P; ++k) s += double(ex_mul(x[w + k], x[w + k]));
                Ec = __double2float_rn(s);
            }
        }
        float pm = warp_incl_scan_maxnan(Ec, lane);
        pm = max_nan(pm, pmcarry);
        pmcarry = __shfl_sync(kFull, pm, 31);
        // running SUM of the window energies (sum form of the drift bound)
        float cs = valid ? Ec * 1.0001f : 0.f;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float t = __shfl_up_sync(kFull, cs, o);
            if (lane >= o) cs += t;
        }
        cs += sumcarry;
        sumcarry = __shfl_sync(kFull, cs, 31);
        uint64_t seed = q;
        float delta = 0.f;
        bool tight = false;
        if (valid) {
            seed = seed_for(c, w, q, avail);
            const bool same = seed == c.a_seed && w >= c.a_w;
            float db = drift_bound(P, same ? w - c.a_w : w - seed, same ? c.a_err : 0.f, pm);
            if (c.sum_ok && same) {
                const float kf = float(w - c.a_w);
                const float growth = kf < 65536.f ? 1.01f : __expf(2.2f * kU * kf) * 1.01f;
                db = fminf(db, kU * (1.5f * float(P) * c.efirst + 2.2f * cs) * growth);
            }
            delta = db + 2.5f * kU * Ec;
            tight = (delta + dscan) <= kTight * Ec;
        }
        const bool bit = (__shfl_sync(kFull, myword, int(r)) >> lane) & 1u;
        const bool possible = Ec >= silent || !(Ec == Ec);
        const bool eval = valid && possible && (bit || !tight);
        uint32_t em = __ballot_sync(kFull, eval);
        if (em == 0) continue;
        float D = 0.f;
        int dec = DEC_NO;
        if (eval) {
            D = dot_avx_order(c, x + w);
            dec = decide_interval(c, D, Ec, delta);
        }
        c.n_cands += __popc(em);
        uint32_t pend = __ballot_sync(kFull, dec != DEC_NO);
        while (pend) {
            const int j = __ffs(pend) - 1;
            pend &= pend - 1;
            const int dj = __shfl_sync(kFull, dec, j);
            const uint64_t wj = wb0 + r * 32 + j;
            const uint64_t sj = __shfl_sync(kFull, seed, j);
            const float Dj = __shfl_sync(kFull, D, j);
            float corr;
            bool hit;
            if (dj == DEC_YES) {
                // the record's corr uses the accurately summed energy (within 1e-5 of the reference's)
                corr = ex_corr(Dj, __shfl_sync(kFull, Ec, j), c.pat->pe);
                hit = true;
            } else {
                hit = exact_window(c, wj, sj, Dj, &corr);
            }
            if (hit) {
                det->c = wj;
                det->seed = sj;
                det->avail = avail_for(c, wj, avail);
                det->corr = corr;
                det->spec = lookup_spec(c, bb, uint32_t(wj - wb0));
                det->status = TMK_ENT_NOSPEC;  // the walk reads the record itself
                det->consumed = 0;
                det->fs = 0;
                det->len = 0;
                return true;
            }
        }
    }
    *pm_out = pmcarry;
    return false;
}

// Fast path of a block (one-shot captures, cluster pass available): when the drift bound at the END of
// the block leaves every unmarked window provably below the threshold ("tight by summary"), only the
// candidate entries the cluster pass prepared (exact dot, accurate energy) have to be decided.
// Returns 0: no detection in the block, 1: detection (*det filled), 2: not applicable (caller falls back
// to process_block).
//   smax      bound on the window energies of the chain up to the end of this block (running maximum)
//   racc_end  bound on the SUM of those window energies, efirst on the energy of the seed window:
//             |E_ref - S| <= u (1.5 P S_seed + 2.2 sum_i S_i) is the same rounding argument as drift_bound()
//             without replacing every term by the maximum, so a loud frame just before the seed does not
//             loosen the bound of the quiet gap behind it
//   emin_lo   lower bound of the window energies of the block's windows >= q
//   pmax_in   bound on the window energies of the chain BEFORE this block (windows >= q)
//   avail     samples fed so far (chunked feeding: the seed of a window depends on it)
__device__ __noinline__ int fast_block(Ctx& c, uint64_t bb, uint64_t q, float4 info, float nb, float smax, float racc_end,
                                       float racc_in, float efirst, float emin_lo, float pmax_in, uint64_t avail,
                                       uint2 res, Detection* det) {
    // res: blk_res of the block (loaded by the caller together with the summaries; meaningful when info.w != 0)
    if (c.blk_res == nullptr || c.P > uint32_t(kFastMaxP) || c.a_w != c.a_seed || c.a_err != 0.f) return 2;
    const float silent = 0.5f * c.thr * c.thr * 1e-6f;
    if (info.y < silent) return 0;  // every window is gated or cannot reach the threshold (as in search())
    const uint64_t wb0 = bb * TMK_BLK;
    uint64_t wlast = wb0 + (TMK_BLK - 1);
    if (wlast >= c.nwin) wlast = c.nwin - 1;
    const uint64_t k_end = wlast - c.a_w;
    const float dscan = kScanErr * nb;
    // the bound of a window of energy S is  d_abs + d_rel * S
    float d_abs, d_rel = 2.5f * kU;
    const uint32_t ptr = info.w != 0.f ? res.x : TMK_ENT_INVALID;
    if (c.C != 0) {
        // Chunked feeding: every process_samples call re-seeds the rolling energy (decoder.rs:133-136), so the chain
        // of a window has at most C + P steps, all of them over windows in [q, w] (a search never starts before q).
        // |E_ref - S| <= u (1.5 P + 2.2 k) max(window energies of the chain) growth: the maximum is bounded by
        // `smax` for the whole block, or window by window by  pmax_in + mgain * S_w  where the cluster pass
        // measured  mgain >= (largest window energy of the block up to w) / S_w.
        uint64_t k = c.C + c.P + TMK_BLK;
        if (k > k_end) k = k_end;
        const float per = drift_bound(c.P, k, 0.f, 1.0f);  // bound per unit of the chain's largest window energy
        d_abs = per * smax;
        if (!(d_abs + dscan <= kTight * emin_lo) || !(d_abs <= 3.0e38f) || !(dscan <= 3.0e38f)) {
            if (info.w == 0.f || !(pmax_in >= 0.f)) return 2;
            if (ptr == TMK_ENT_INVALID) return 2;
            const float mgain = __uint_as_float(res.y);
            const float hist = per * pmax_in;
            const float rel = per * mgain + 2.5f * kU;
            if (!(hist + dscan <= 0.5f * kTight * emin_lo) || !(rel <= 0.5f * kTight)) return 2;
            d_abs = hist;
            d_rel = rel;
        }
    } else {
    const float dmax = drift_bound(c.P, k_end, 0.f, smax);
    const float kf = float(k_end);
    const float growth = kf < 65536.f ? 1.01f : __expf(2.2f * kU * kf) * 1.01f;
    const float dsum = kU * (1.5f * float(c.P) * efirst + 2.2f * racc_end) * growth;
    const float d_end = fminf(dmax, dsum);  // both are rigorous; NaN in either falls through below
    d_abs = d_end;
    if (!(d_end + dscan <= kTight * emin_lo) || !(d_end <= 3.0e38f) || !(dscan <= 3.0e38f)) {
        // A block whose energy RISES (the quiet gap runs into a loud preamble) fails the test above only
        // because the drift at the END of the block is held against the energy at its START. Window by
        // window the sum form is  u (1.5 P S_seed + 2.2 (racc_in + B_w)),  B_w = the window energies of
        // this block up to w; the cluster pass measured  gain >= B_w / S_w  for every window of the block.
        if (info.w == 0.f || !(racc_in >= 0.f)) return 2;
        if (ptr == TMK_ENT_INVALID) return 2;
        const float gain = __uint_as_float(res.y);
        const float hist = kU * (1.5f * float(c.P) * efirst + 2.2f * racc_in) * growth;
        const float rel = 2.2f * kU * growth * gain + 2.5f * kU;
        if (!(hist + dscan <= 0.5f * kTight * emin_lo) || !(rel <= 0.5f * kTight)) return 2;
        d_abs = hist;
        d_rel = rel;
    }
    }
    if (info.w == 0.f) return 0;
    if (ptr == TMK_ENT_INVALID) return 2;
    const uint32_t n = ptr & 63u;
    for (uint32_t r0 = 0; r0 < n; r0 += 32) {
        TmkCandEntry e;
        e.wl = 0; e.status = TMK_ENT_NOSPEC; e.D = 0.f; e.Ec = 0.f; e.spec = TMK_ENT_INVALID; e.consumed = 0; e.fs_off = 0; e.len = 0;
        if (r0 + uint32_t(c.lane) < n) e = ld_entry(c.entries + (ptr >> 6) + r0 + c.lane);
        const uint64_t w = wb0 + e.wl;
        const bool valid = r0 + uint32_t(c.lane) < n && w >= q && w < c.nwin;
        int dec = DEC_NO;
        if (valid) dec = decide_interval(c, e.D, e.Ec, d_abs + d_rel * e.Ec);
        c.n_cands += __popc(__ballot_sync(kFull, valid));
        uint32_t pend = __ballot_sync(kFull, dec != DEC_NO);
        while (pend) {
            const int j = __ffs(pend) - 1;
            pend &= pend - 1;
            const int dj = __shfl_sync(kFull, dec, j);
            const uint64_t wj = __shfl_sync(kFull, w, j);
            const float Dj = __shfl_sync(kFull, e.D, j);
            const float Ecj = __shfl_sync(kFull, e.Ec, j);
            const uint32_t sj = __shfl_sync(kFull, e.spec, j);
            const uint64_t seedj = seed_for(c, wj, q, avail);  // == q for one-shot captures
            float corr;
            bool hit;
            if (dj == DEC_YES) {
                corr = ex_corr(Dj, Ecj, c.pat->pe);
                hit = true;
            } else {
                hit = exact_window(c, wj, seedj, Dj, &corr);
            }
            if (hit) {
                det->c = wj;
                det->seed = seedj;
                det->avail = avail_for(c, wj, avail);
                det->corr = corr;
                det->spec = sj;
                det->status = __shfl_sync(kFull, uint32_t(e.status), j);
                det->consumed = __shfl_sync(kFull, e.consumed, j);
                det->fs = wj + int64_t(__shfl_sync(kFull, e.fs_off, j));
                det->len = __shfl_sync(kFull, e.len, j);
                return 1;
            }
        }
    }
    return 0;
}

// search_for_preamble over all future calls: first window >= q with corr_ref >= thr.
__device__ __noinline__ bool search(Ctx& c, uint64_t q, uint64_t avail, Detection* det) {
    if (c.nwin == 0 || q >= c.nwin) return false;
    const int lane = c.lane;
    const uint32_t P = c.P;
    const float silent = 0.5f * c.thr * c.thr * 1e-6f;
    c.a_seed = q;
    c.a_w = q;
    c.a_err = 0.f;
    float runmax = 0.f;  // bound on the window energies in (a_w, first window of the group)
    // the block that contains q is partial: its summary also covers windows before q (the frame that was
    // just consumed). The fast path bounds the drift by the SUM of the energies, which that frame does not
    // enter; otherwise the block is examined window by window.
    uint64_t b0 = q / TMK_BLK;
    float racc = 0.f, efirst = 0.f;  // sum / seed-window bounds of the chain (see fast_block)
    // the summaries and cluster results of the first group of blocks behind the one that holds q start their trip
    // now, together with those of that block: one memory round trip instead of two
    const uint64_t pre_b = b0 + 1;
    float4 pre_info = make_float4(0.f, 0.f, 0.f, 0.f);
    uint2 pre_res = make_uint2(TMK_ENT_INVALID, 0u);
    if (pre_b + lane < c.nblk) {
        pre_info = c.blk_info[pre_b + lane];
        if (c.blk_res != nullptr) pre_res = __ldcg(c.blk_res + pre_b + lane);
    }
    {
        const float4 i0 = c.blk_info[b0];
        const uint2 res0 = c.blk_res != nullptr ? __ldcg(c.blk_res + b0) : make_uint2(TMK_ENT_INVALID, 0u);
        float nb0 = i0.y;
        if (b0 > 0) nb0 = max_nan(nb0, c.blk_info[b0 - 1].y);
        if (b0 + 1 < c.nblk) nb0 = max_nan(nb0, c.blk_info[b0 + 1].y);
        uint64_t wl0 = b0 * TMK_BLK + (TMK_BLK - 1);
        if (wl0 >= c.nwin) wl0 = c.nwin - 1;
        const float dscan0 = kScanErr * nb0;
        // 1. summary only: every window >= q bounded by the block maximum (which the consumed frame may own)
        efirst = i0.y * 1.001f;
        racc = float(wl0 - q + 1) * i0.y * 1.001f;
        // (a block with a strong contrast is the end of the consumed frame: its maximum would pollute the
        // bounds of the whole gap behind it, so it goes straight to the exact energies)
        int r = 2;
        if (!(i0.y > float(TMK_CONTRAST) * i0.x)) {
            r = fast_block(c, b0, q, i0, nb0, i0.y * 1.001f, racc, 0.f, efirst, i0.x - dscan0, 0.f, avail, res0, det);
            if (r == 1) return true;
            if (r == 0) runmax = i0.y * 1.001f;
        }
        if (r == 2 && c.blk_res != nullptr && c.P <= uint32_t(kFastMaxP)) {
            // 2. exact energies of the windows [q, end of block] (one block load): sum, maximum, minimum
            block_prefix(c, b0);
            const uint32_t lo = uint32_t(q - b0 * TMK_BLK), hi = uint32_t(wl0 - b0 * TMK_BLK);
            float es = 0.f, emx = 0.f, emn = __int_as_float(0x7f800000), eq = 0.f;
            for (uint32_t wl = lo + lane; wl <= hi; wl += 32) {
                const float ec = __double2float_rn(c.sc->Q[wl + P] - c.sc->Q[wl]);
                es += ec;
                emx = max_nan(emx, ec);
                emn = fminf(emn, ec);
                if (wl == lo) eq = ec;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                es += __shfl_xor_sync(kFull, es, o);
                emx = max_nan(emx, __shfl_xor_sync(kFull, emx, o));
                emn = fminf(emn, __shfl_xor_sync(kFull, emn, o));
            }
            eq = __shfl_sync(kFull, eq, 0);
            const float racc_x = es * 1.0001f;
            r = fast_block(c, b0, q, i0, nb0, emx * 1.0001f, racc_x, 0.f, eq * 1.0001f, emn * 0.9999f, 0.f, avail, res0, det);
            if (r == 1) return true;
            if (r == 0) {
                runmax = emx * 1.0001f;
                racc = racc_x;
                efirst = eq * 1.0001f;
            }
        }
        c.sum_ok = c.C == 0;
        c.efirst = efirst;
        c.racc_in = 0.f;
        if (r == 2 && process_block(c, b0, q, avail, 0.f, det, &runmax)) return true;
    }
    ++b0;
    while (b0 < c.nblk) {
        c.dbg[3]++;
        const bool sum_ok = c.C == 0 && c.a_w == c.a_seed && c.a_err == 0.f;
        const uint64_t bb = b0 + lane;
        const bool valid = bb < c.nblk;
        float4 info = make_float4(0.f, 0.f, 0.f, 0.f);
        uint2 res = make_uint2(TMK_ENT_INVALID, 0u);
        bool any = false;
        if (b0 == pre_b) {  // the group that was requested at the entry of the search
            info = pre_info;
            res = pre_res;
            any = valid && info.w != 0.f;
        } else if (valid) {
            info = c.blk_info[bb];
            if (c.blk_res != nullptr) res = __ldcg(c.blk_res + bb);
            any = info.w != 0.f;
        }
        // neighbourhood maximum (blocks bb-1 .. bb+1) for the scan kernel's own approximation error
        float nb = max_nan(info.y, max_nan(__shfl_up_sync(kFull, info.y, 1), __shfl_down_sync(kFull, info.y, 1)));
        if (valid && lane == 0 && bb > 0) nb = max_nan(nb, c.blk_info[bb - 1].y);
        if (valid && lane == 31 && bb + 1 < c.nblk) nb = max_nan(nb, c.blk_info[bb + 1].y);
        const float incl = warp_incl_scan_maxnan(info.y, lane);
        const float up = __shfl_up_sync(kFull, incl, 1);
        const float excl = lane == 0 ? runmax : max_nan(runmax, up);
        const float run = max_nan(runmax, incl);
        bool need = false, hist_loose = false;
        uint64_t wfirst = bb * TMK_BLK, s_first = q, kk = 0;
        // sum of the window energies block by block (exclusive prefix over the lanes), for fast_block
        float contrib = 0.f;
        if (valid) {
            uint64_t wl2 = wfirst + (TMK_BLK - 1);
            if (wl2 >= c.nwin) wl2 = c.nwin - 1;
            contrib = float(wl2 - wfirst + 1) * info.y * 1.001f;
        }
        float cincl = contrib;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float t = __shfl_up_sync(kFull, cincl, o);
            if (lane >= o) cincl += t;
        }
        if (valid) {
            uint64_t wlast = wfirst + (TMK_BLK - 1);
            if (wlast >= c.nwin) wlast = c.nwin - 1;
            s_first = seed_for(c, wfirst, q, avail);  // earliest seed of any window in the block
            const bool same = s_first == c.a_seed;
            kk = wlast - (same ? c.a_w : s_first);
            const float e0 = same ? c.a_err : 0.f;
            float dbi = drift_bound(P, kk, e0, run * 1.001f), dbe = drift_bound(P, kk, e0, excl * 1.001f);
            if (sum_ok && same) {
                const float kf = float(kk);
                const float growth = kf < 65536.f ? 1.01f : __expf(2.2f * kU * kf) * 1.01f;
                dbi = fminf(dbi, kU * (1.5f * float(P) * efirst + 2.2f * (racc + cincl)) * growth);
                dbe = fminf(dbe, kU * (1.5f * float(P) * efirst + 2.2f * (racc + cincl - contrib)) * growth);
            }
            const float delta = dbi + kScanErr * nb + 2.5f * kU * info.y;
            // (1) no window of the block can reach the threshold even with the lowest possible E_ref:
            //     corr_ref^2 = corr_true^2 * S / E_ref  with  E_ref >= max(S - delta, 1e-6)
            float gain;  // upper bound of S / E_ref over the block
            if (delta < 0.5f * info.x) gain = 1.0f / (1.0f - delta / (info.x * 0.999f));
            else gain = info.y * 1.001f * 1e6f;
            const bool fine = info.x >= 1e-3f * nb;  // the scan's corr^2 is accurate for every window
            const bool safe = fine && info.z * 1.003f * gain < c.thr * c.thr * 0.999f;
            const bool is_silent = info.y < silent;
            need = !is_silent && (any || !safe);
            // (2) is the drift accumulated BEFORE this block what makes its windows undecidable?
            hist_loose = dbe > 0.25f * kTight * info.x;
        }
        const uint32_t m = __ballot_sync(kFull, need);
        if (m == 0) {
            runmax = __shfl_sync(kFull, run, 31);
            racc += __shfl_sync(kFull, cincl, 31);
            b0 += 32;
            continue;
        }
        const int j = __ffs(m) - 1;
        {
            const float pmax_j = __shfl_sync(kFull, excl, j);
            float4 ij;
            ij.x = __shfl_sync(kFull, info.x, j);
            ij.y = __shfl_sync(kFull, info.y, j);
            ij.z = __shfl_sync(kFull, info.z, j);
            ij.w = __shfl_sync(kFull, info.w, j);
            const float nbj = __shfl_sync(kFull, nb, j);
            const float racc_end_j = racc + __shfl_sync(kFull, cincl, j);
            const float racc_in_j = racc_end_j - __shfl_sync(kFull, contrib, j);
            uint2 rj;
            rj.x = __shfl_sync(kFull, res.x, j);
            rj.y = __shfl_sync(kFull, res.y, j);
            const int r = fast_block(c, b0 + j, q, ij, nbj, max_nan(pmax_j, ij.y * 1.001f), racc_end_j, racc_in_j, efirst,
                                     ij.x - kScanErr * nbj, pmax_j, avail, rj, det);
            if (r == 1) return true;
            if (r == 0) {
                runmax = max_nan(pmax_j, ij.y * 1.001f);
                racc = racc + __shfl_sync(kFull, cincl, j);
                b0 += j + 1;
                continue;
            }
        }
        racc += __shfl_sync(kFull, cincl, j);  // the slow paths below continue behind block j (or re-anchor at it)
        const uint64_t wf = __shfl_sync(kFull, wfirst, j), sf = __shfl_sync(kFull, s_first, j);
        const uint64_t kj = __shfl_sync(kFull, kk, j);
        const bool hl = __shfl_sync(kFull, int(hist_loose), j) != 0;
        const bool one_seed = c.C == 0 || seed_for(c, wf + TMK_BLK - 1 < c.nwin ? wf + TMK_BLK - 1 : c.nwin - 1, q, avail) == sf;
        if (hl && one_seed && kj > 1024) {
            // re-anchor: roll the reference's chain up to this block once and measure its drift
            c.dbg[1]++;
            const float eref = exact_energy(c, sf, wf);
            const float ec = window_energy(c, wf);
            float err = fabsf(eref - ec) + 1.2f * kU * ec;
            if (!(err == err)) err = __int_as_float(0x7f800000);
            c.a_seed = sf;
            c.a_w = wf;
            c.a_err = err;
            runmax = 0.f;
            b0 += j;
            continue;  // re-evaluate from this block with the measured anchor
        }
        const float pmax = __shfl_sync(kFull, excl, j);
        c.sum_ok = sum_ok;
        c.efirst = efirst;
        c.racc_in = racc - __shfl_sync(kFull, contrib, j);  // racc already includes block j here
        float pm_blk;
        if (process_block(c, b0 + j, q, avail, pmax, det, &pm_blk)) return true;
        // continue behind block j; the running maximum now covers it (pm_blk: exact energies of the
        // block's windows >= q on top of pmax)
        runmax = pm_blk;
        b0 += j + 1;
    }
    return false;
}

// ---------------------------------------------------------------- refinement (decoder.rs:157-221)
__device__ __noinline__ uint64_t refine(const Ctx& c, const Detection& d) {
    const TmkPattern* pat = c.pat;
    const uint32_t P = c.P, sl = pat->sync_len, margin = pat->margin;
    const float* __restrict__ x = c.x;
    const uint64_t expect = d.c + P - sl;
    const uint64_t e_rel = expect - d.seed;  // relative to the search area start
    const uint64_t lo = d.seed + (e_rel > margin ? e_rel - margin : 0);
    uint64_t hi = expect + margin;
    if (hi > d.avail - sl) hi = d.avail - sl;
    float best = -1.0f;
    uint64_t best_off = expect;
    // stage the samples [lo, hi + sl) in shared memory with one coalesced round of loads (the prefix-sum
    // buffer of process_block is dead here); every lane then walks its own offset out of shared memory
    const uint64_t span = hi >= lo ? hi + sl - lo : 0;
    float* xs = reinterpret_cast<float*>(c.sc->Q);
    const bool staged = span <= 2 * kQBuf;
    if (staged) {
        __syncwarp();
        for (uint32_t i = c.lane; i < uint32_t(span); i += 32) xs[i] = x[lo + i];
        __syncwarp();
    }
    for (uint64_t base = lo; base <= hi; base += 32) {
        const uint64_t j = base + c.lane;
        float corr = -2.0f;
        if (j <= hi) {
            float dot = 0.0f, we = 0.0f;
            const float* wp = staged ? xs + (j - lo) : x + j;
            for (uint32_t t = 0; t < sl; ++t) {
                const float w = wp[t];
                dot = ex_add(dot, mul_pm1(w, c.sc->pre[P - sl + t]));
                we = ex_add(we, ex_mul(w, w));
            }
            corr = (we > 1e-6f && pat->sync_energy > 1e-6f) ? ex_div(dot, ex_mul(ex_sqrt(we), pat->sync_energy)) : 0.0f;
            if (!(corr == corr)) corr = -2.0f;  // NaN never satisfies `corr > best_corr`
        }
        // first strict maximum: larger value wins, ties go to the lower offset
        float v = corr;
        int idx = c.lane;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(kFull, v, o);
            const int oi = __shfl_xor_sync(kFull, idx, o);
            if (ov > v || (ov == v && oi < idx)) {
                v = ov;
                idx = oi;
            }
        }
        if (v > best) {
            best = v;
            best_off = base + idx;
        }
    }
    return best_off + sl;
}

// ---------------------------------------------------------------- line decode
// decode_4b5b_symbol, line_coding.rs:157-180, as two packed tables: nibble of symbol s at bits 4 (s % 8) of
// kDec5Nib[s / 8], validity at bit s of kDec5Valid
constexpr int8_t kDec5[32] = {-1, -1, -1, -1, -1, -1, -1, -1, -1, 1, 4, 5, -1, -1, 6, 7,
                              -1, -1, 8, 9, 2, 3, 10, 11, -1, -1, 12, 13, 14, 15, 0, -1};
constexpr uint32_t dec5_nib_word(int w) {
    uint32_t v = 0;
    for (int k = 0; k < 8; ++k) v |= uint32_t(kDec5[8 * w + k] < 0 ? 0 : kDec5[8 * w + k]) << (4 * k);
    return v;
}
constexpr uint32_t dec5_valid() {
    uint32_t v = 0;
    for (int k = 0; k < 32; ++k) v |= uint32_t(kDec5[k] >= 0) << k;
    return v;
}
constexpr uint32_t kDec5Nib1 = dec5_nib_word(1), kDec5Nib2 = dec5_nib_word(2), kDec5Nib3 = dec5_nib_word(3);
constexpr uint32_t kDec5Valid = dec5_valid();
static_assert(dec5_nib_word(0) == 0 && kDec5Valid == 0x7CFCCE00u, "4B5B decode table");
__device__ __forceinline__ int dec5(uint32_t sym) {  // nibble or -1
    const uint32_t w = (sym >> 3) == 1 ? kDec5Nib1 : (sym >> 3) == 2 ? kDec5Nib2 : kDec5Nib3;
    return ((kDec5Valid >> sym) & 1u) ? int((w >> (4 * (sym & 7))) & 15u) : -1;
}

// f32 SUM of one level: sequential from 0 like the reference's mean (line_coding.rs:104-113, 258-261); the
// division by spl is not performed here (see line_decode)
__device__ __forceinline__ float level_sum(const float* __restrict__ s, uint32_t spl) {
    float acc = 0.0f;
    for (uint32_t i = 0; i < spl; ++i) acc = ex_add(acc, s[i]);
    return acc;
}

// LineCode::decode over the samples of `total_bits` bits starting at fs. Writes MSB-first packed
// bytes (crc.rs:63-75) to c.sc->bytes and returns the number of bits decoded (4B5B stops at the
// first invalid symbol, line_coding.rs:281-286).
//
// The reference decides on level MEANS avg = fl(sum / spl). The decisions only need
//   Manchester (:115-119)   avg_a > avg_b
//   4B5B + NRZI (:262-270)  last_avg * avg < 0   and   |avg| > 1e-6
// and x -> fl(x / spl) is monotone, so they are taken on the level SUMS whenever that is provably the same:
//   |avg| > 1e-6  <=>  |sum| >= lev_thr (the smallest sum whose quotient exceeds 1e-6, found on the host);
//   the sign of last_avg * avg is the product of the signs unless the product can underflow to +-0, i.e. unless
//   |sum| < 1e-30 (|last_avg| > 1e-6 always) — NaN sums fail the same test;
//   sum_a > sum_b with a relative gap above 1e-6 (8 ulps) keeps avg_a > avg_b after both roundings.
// Everything else (digital silence, NaN, near ties) takes the literal path with the IEEE divisions. The sums drop
// the reference's leading `0.0 +`, which can only turn a -0.0 sum into +0.0: a zero mean decides the same either way.
__device__ __noinline__ uint32_t line_decode(const Ctx& c, uint64_t fs, uint32_t total_bits) {
    const float* __restrict__ s = c.x + fs;
    const uint32_t spl = c.pat->spl;
    const float dv = float(spl), lev_thr = c.pat->lev_thr;
    const int lane = c.lane;
    uint8_t* bytes = c.sc->bytes;
    // level sums are staged in shared memory (the prefix-sum buffer is idle during a decode):
    // pass 1 fills a chunk with independent loads in flight, pass 2 decides out of shared memory
    float* lev = reinterpret_cast<float*>(c.sc->Q);
    constexpr uint32_t kChunk = 640;  // levels per chunk: 128 4B5B symbols = 64 bytes, 320 Manchester bits = 40 bytes
    static_assert(kChunk <= 2 * kQBuf && kChunk % 160 == 0 && kChunk % 64 == 0, "level chunk");
    const bool man = c.pat->line_code == TMK_LINE_MANCHESTER;
    const uint32_t nlev = man ? total_bits * 2 : (total_bits / 4) * 5;
    float last = dv;  // prev_level_avg restarts at 1.0 on every decode call (line_coding.rs:251): the SUM whose mean is 1.0
    uint32_t nb = 0;
    for (uint32_t c0 = 0; c0 < nlev; c0 += kChunk) {
        const uint32_t nl = nlev - c0 < kChunk ? nlev - c0 : kChunk;
        __syncwarp();
        if (spl == 3) {
            // 8 rounds of loads (24 per lane) in flight before the first store: the stores to `lev` may alias
            // the samples as far as the compiler knows, so a plain loop pays one HBM round trip per round
            constexpr int kB = 8;
            for (uint32_t i0 = 0; i0 < nl; i0 += 32 * kB) {
                float a[kB][3];
#pragma unroll
                for (int u = 0; u < kB; ++u) {
                    const uint32_t i = i0 + u * 32 + lane;
                    if (i < nl) {
                        const float* p = s + uint64_t(c0 + i) * 3;
                        a[u][0] = p[0];
                        a[u][1] = p[1];
                        a[u][2] = p[2];
                    }
                }
#pragma unroll
                for (int u = 0; u < kB; ++u) {
                    const uint32_t i = i0 + u * 32 + lane;
                    if (i < nl) lev[i] = ex_add(ex_add(a[u][0], a[u][1]), a[u][2]);
                }
            }
        } else {
#pragma unroll 2
            for (uint32_t i = lane; i < nl; i += 32) lev[i] = level_sum(s + uint64_t(c0 + i) * spl, spl);
        }
        __syncwarp();
        if (man) {
            // one lane per bit: first half > second half -> 0, else 1 (tie -> 1)  line_coding.rs:104-119
            const uint32_t nbit = nl / 2;
#pragma unroll 1
            for (uint32_t b0 = 0; b0 < nbit; b0 += 32) {
                const uint32_t bi = b0 + lane;
                bool bit = false;
                if (bi < nbit) {
                    const float2 ab = *reinterpret_cast<const float2*>(lev + 2 * bi);
                    const float m = fmaxf(fabsf(ab.x), fabsf(ab.y));
                    if (fabsf(ab.x - ab.y) > 1e-6f * m && m >= 1e-30f && m <= 1e37f) bit = !(ab.x > ab.y);
                    else bit = !(ex_div(ab.x, dv) > ex_div(ab.y, dv));
                }
                const uint32_t rev = __brev(__ballot_sync(kFull, bit));
                if (lane < 4 && b0 + lane * 8 < nbit) bytes[(nb >> 3) + lane] = uint8_t(rev >> (24 - 8 * lane));
                nb += nbit - b0 < 32 ? nbit - b0 : 32;
            }
            continue;
        }
        // 4B5B + NRZI: one lane per 5-level symbol, 32 symbols (16 bytes) per step
        const uint32_t nsym = nl / 5;
        bool stop = false;
#pragma unroll 1
        for (uint32_t s0 = 0; s0 < nsym && !stop; s0 += 32) {
            const uint32_t si = s0 + lane;
            const bool active = si < nsym;
            float v[5];
#pragma unroll
            for (int j = 0; j < 5; ++j) v[j] = active ? lev[5 * si + j] : 1.0f;
            bool plain = true;  // every sum of the symbol is far from the underflow of last_avg * avg (and not NaN)
#pragma unroll
            for (int j = 0; j < 5; ++j) plain = plain && fabsf(v[j]) >= 1e-30f;
            uint32_t sym = 0, hm;
            float mylast = 0.0f;
            if (__all_sync(kFull, plain)) {
                // the last level of this symbol with |avg| > 1e-6, if any (line_coding.rs:268-270)
                bool have = false;
#pragma unroll
                for (int j = 0; j < 5; ++j)
                    if (fabsf(v[j]) >= lev_thr) {
                        mylast = v[j];
                        have = true;
                    }
                hm = __ballot_sync(kFull, have && active);
                const uint32_t below = hm & ((1u << lane) - 1u);
                const float pv = __shfl_sync(kFull, mylast, below ? 31 - __clz(below) : 0);
                uint32_t lastbits = __float_as_uint(below ? pv : last);
#pragma unroll
                for (int j = 0; j < 5; ++j) {
                    const uint32_t vb = __float_as_uint(v[j]);
                    sym = (sym << 1) | ((lastbits ^ vb) >> 31);  // :262-266, first level = MSB
                    if (fabsf(v[j]) >= lev_thr) lastbits = vb;
                }
            } else {
                // literal: the means, their products, the 1e-6 gate
                float av[5];
                bool have = false;
#pragma unroll
                for (int j = 0; j < 5; ++j) {
                    av[j] = ex_div(v[j], dv);
                    if (fabsf(av[j]) > 1e-6f) {
                        mylast = v[j];
                        have = true;
                    }
                }
                hm = __ballot_sync(kFull, have && active);
                const uint32_t below = hm & ((1u << lane) - 1u);
                const float pv = __shfl_sync(kFull, mylast, below ? 31 - __clz(below) : 0);
                float lastv = ex_div(below ? pv : last, dv);
#pragma unroll
                for (int j = 0; j < 5; ++j) {
                    sym = (sym << 1) | (ex_mul(lastv, av[j]) < 0.0f ? 1u : 0u);
                    if (fabsf(av[j]) > 1e-6f) lastv = av[j];
                }
            }
            if (hm) last = __shfl_sync(kFull, mylast, 31 - __clz(hm));
            const int nib = dec5(sym);
            const uint32_t nsy = nsym - s0 < 32 ? nsym - s0 : 32;
            const uint32_t inv = __ballot_sync(kFull, active && nib < 0);
            const uint32_t nval = inv ? uint32_t(__ffs(inv) - 1) : nsy;  // stop at the first invalid symbol :281-286
            const int lo = __shfl_down_sync(kFull, nib, 1);
            if ((lane & 1) == 0 && uint32_t(lane) < nval)
                bytes[(nb >> 3) + (lane >> 1)] = uint8_t((uint32_t(nib) << 4) | (uint32_t(lane) + 1 < nval ? uint32_t(lo) : 0u));
            nb += 4 * nval;
            if (inv) stop = true;
        }
        if (stop) break;
    }
    __syncwarp();
    return nb;
}

// CRC-8 (poly 0x07, init 0; crc.rs:7-22) of bytes[0..n), lane-parallel:
// crc = XOR_i  b_i(x) * x^(8(n-i)) mod g
__device__ uint32_t crc8_warp(const uint8_t* bytes, uint32_t n, int lane) {
    uint32_t acc = 0;
    for (uint32_t i = lane; i < n; i += 32) {
        const uint32_t a = bytes[i], w = g_crc_pow[n - i];
        uint32_t p = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) p ^= ((a >> j) & 1u) ? (w << j) : 0u;
#pragma unroll
        for (int k = 14; k >= 8; --k) p ^= ((p >> k) & 1u) ? (0x107u << (k - 8)) : 0u;
        acc ^= p & 0xffu;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc ^= __shfl_xor_sync(kFull, acc, o);
    return acc;
}

// ---------------------------------------------------------------- verdict records
struct Emit {
    tmk_frame_rec* recs;
    uint4* rec_ord;  // per record slot: (speculative record to gather or TMK_ENT_INVALID, order in the stream, stream, corr)
    uint64_t rec_cap;
    uint8_t* payload;
    uint64_t payload_cap;
    TmkCounters* counters;
    uint32_t stream;
    uint32_t ord;  // records emitted by this stream so far
    // record slots are reserved 32 at a time per warp (one atomic per 32 records instead of one per record); the
    // slots a warp leaves unused are marked in rec_ord (kNoRecord) and skipped by the ordering kernel
    unsigned long long slot_cur, slot_end;
};
constexpr uint32_t kNoRecord = 0xffffffffu;

__device__ __forceinline__ unsigned long long take_slot(int lane, Emit& em) {
    if (em.slot_cur == em.slot_end) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&em.counters->n_recs, 32ull);
        base = __shfl_sync(kFull, base, 0);
        em.slot_cur = base;
        em.slot_end = base + 32;
    }
    return em.slot_cur++;
}

__device__ __noinline__ void emit(const Ctx& c, Emit& em, uint32_t status, const Detection& d, uint64_t frame_start,
                     uint32_t consumed, bool have_hdr, const uint8_t* hdr, uint32_t data_len) {
    const unsigned long long slot = take_slot(c.lane, em);
    unsigned long long poff = 0;
    if (c.lane == 0 && data_len) poff = atomicAdd(&em.counters->n_payload, (unsigned long long)data_len);
    poff = __shfl_sync(kFull, poff, 0);
    if (data_len && poff + data_len <= em.payload_cap) {
        for (uint32_t i = c.lane; i < data_len; i += 32) em.payload[poff + i] = c.sc->bytes[TMK_PHY_HEADER_BYTES + i];
    }
    if (c.lane == 0 && slot < em.rec_cap) {
        tmk_frame_rec r;
        r.stream = em.stream;
        r.status = status;
        r.preamble_start = d.c;
        r.frame_start = frame_start;
        r.payload_off = poff;
        r.corr = d.corr;
        r.consumed = consumed;
        r.len = have_hdr ? uint16_t((uint32_t(hdr[0]) << 8) | hdr[1]) : 0;
        r.crc = have_hdr ? hdr[2] : 0;
        r.type = have_hdr ? hdr[3] : 0;
        r.seq = have_hdr ? hdr[4] : 0;
        r.src = have_hdr ? hdr[5] : 0;
        r.dst = have_hdr ? hdr[6] : 0;
        r.hdr_valid = have_hdr ? 1 : 0;
        em.recs[slot] = r;
        em.rec_ord[slot] = make_uint4(TMK_ENT_INVALID, em.ord, em.stream, __float_as_uint(d.corr));
    }
    em.ord++;
}

__device__ __forceinline__ void init_ctx(Ctx& c, const float* samples, const TmkStreamDev& sd, const TmkPattern* patterns,
                                         const float4* blk_info, const uint32_t* cand_bits, const TmkFast& f, int lane,
                                         uint32_t warp) {
    c.x = samples + sd.offset;
    c.N = sd.n_samples;
    c.pat = patterns + sd.pattern;
    c.P = c.pat->P;
    c.nwin = c.N >= c.P ? c.N - c.P + 1 : 0;
    c.C = (sd.chunk == 0 || sd.chunk >= c.N) ? 0 : sd.chunk;
    c.blk_info = blk_info + sd.blk_base;
    c.cand_bits = cand_bits + sd.blk_base * TMK_BLK_WORDS;
    c.blk_res = f.blk_res ? f.blk_res + sd.blk_base : nullptr;
    c.entries = f.entries;
    c.spec = f.spec;
    c.nblk = (c.nwin + TMK_BLK - 1) / TMK_BLK;
    c.thr = sd.threshold;
    c.local_addr = sd.local_addr;
    c.max_frame_bytes = sd.max_frame_bytes;
    c.lane = lane;
    c.sc.w = warp;
    const ScratchRef sc = c.sc;
    c.rvalid = false;
    c.rseed = c.rpos = 0;
    c.rE = 0.f;
    c.a_seed = c.a_w = 0;
    c.a_err = 0.f;
    c.racc_in = c.efirst = 0.f;
    c.sum_ok = false;
    c.n_replays = c.n_cands = c.n_det = 0;
    c.dbg[0] = c.dbg[1] = c.dbg[2] = c.dbg[3] = 0;
    c.n_par_steps = 0;
    c.n_par_doubt = 0;
    __syncwarp();
    for (uint32_t k = lane; k < c.P; k += 32) sc->pre[k] = int8_t(c.pat->levels[k / c.pat->spl]);
    __syncwarp();
}

// ---------------------------------------------------------------- cluster pass
// decode_frame (decoder.rs:242-356) for a detection at window w of a ONE-SHOT capture (the whole capture is
// buffered: avail == N), exactly the sequence of the walk kernel below. Fills a verdict record; the payload
// goes straight into the payload buffer.
__device__ __noinline__ void spec_decode(Ctx& c, uint64_t w, float corr, uint32_t stream, tmk_frame_rec* out,
                                         uint8_t* payload, uint64_t payload_cap, TmkCounters* counters) {
    Detection d;
    d.c = w;
    d.seed = 0;  // refinement range [expect - margin, expect + margin]: the walk checks that its seed agrees
    d.avail = c.N;
    d.corr = corr;
    d.spec = TMK_ENT_INVALID;
    const int lc = c.pat->line_code;
    const uint32_t spl = c.pat->spl, hs = c.pat->hdr_samples;
    const uint64_t fs = refine(c, d);
    uint32_t status, consumed = 0, len = 0, data_len = 0;
    bool have_hdr = false;
    uint8_t hdr[TMK_PHY_HEADER_BYTES];
    for (int i = 0; i < TMK_PHY_HEADER_BYTES; ++i) hdr[i] = 0;
    if (c.N < fs + hs) {
        status = TMK_FRAME_TRUNCATED;
    } else {
        const uint8_t* hb = c.sc->bytes;
        const uint32_t hbits = line_decode(c, fs, 8 * TMK_PHY_HEADER_BYTES);
        const bool have7 = (hbits + 7) / 8 >= TMK_PHY_HEADER_BYTES;
        for (int i = 0; i < TMK_PHY_HEADER_BYTES; ++i) hdr[i] = have7 ? hb[i] : 0;
        __syncwarp();
        const uint32_t type = hdr[3];
        len = (uint32_t(hdr[0]) << 8) | hdr[1];
        if (!have7 || (type != TMK_FRAME_TYPE_DATA && type != TMK_FRAME_TYPE_ACK)) {
            status = TMK_FRAME_HDR_PARSE;
            consumed = hs;
            have_hdr = have7;
        } else if ((type == TMK_FRAME_TYPE_DATA && len == 0) || len > c.max_frame_bytes) {
            status = TMK_FRAME_BAD_LEN;
            consumed = 1;
            have_hdr = true;
        } else {
            const uint32_t total_bits = (TMK_PHY_HEADER_BYTES + len) * 8;
            const uint64_t total_samples = tmk_sfb(lc, spl, total_bits);
            if (c.N < fs + total_samples) {
                status = TMK_FRAME_TRUNCATED;
            } else {
                const uint32_t nbits = line_decode(c, fs, total_bits);
                consumed = c.P + uint32_t(tmk_sfb(lc, spl, nbits));
                have_hdr = true;
                if (nbits < total_bits) {
                    status = TMK_FRAME_SHORT_DECODE;
                } else if (c.local_addr >= 0 && hdr[6] != uint32_t(c.local_addr)) {
                    status = TMK_FRAME_NOT_FOR_US;
                    data_len = len;
                } else {
                    const uint32_t crc = crc8_warp(c.sc->bytes + TMK_PHY_HEADER_BYTES, len, c.lane);
                    status = crc != hdr[2] ? TMK_FRAME_CRC_FAIL : TMK_FRAME_OK;
                    data_len = len;
                }
            }
        }
    }
    unsigned long long poff = 0;
    if (c.lane == 0 && data_len) poff = atomicAdd(&counters->n_payload, (unsigned long long)data_len);
    poff = __shfl_sync(kFull, poff, 0);
    if (data_len && poff + data_len <= payload_cap)
        for (uint32_t i = c.lane; i < data_len; i += 32) payload[poff + i] = c.sc->bytes[TMK_PHY_HEADER_BYTES + i];
    if (c.lane == 0) {
        tmk_frame_rec r;
        r.stream = stream;
        r.status = status;
        r.preamble_start = w;
        r.frame_start = fs;
        r.payload_off = poff;
        r.corr = corr;
        r.consumed = consumed;
        r.len = have_hdr ? uint16_t(len) : 0;
        r.crc = have_hdr ? hdr[2] : 0;
        r.type = have_hdr ? hdr[3] : 0;
        r.seq = have_hdr ? hdr[4] : 0;
        r.src = have_hdr ? hdr[5] : 0;
        r.dst = have_hdr ? hdr[6] : 0;
        r.hdr_valid = have_hdr ? 1 : 0;
        *out = r;
    }
    __syncwarp();
}

// result word of a block: the walk of the chunk starts when this kernel has finished, so plain stores suffice
__device__ __forceinline__ void publish_blk(uint2* p, uint32_t v, float gain) {
    *p = make_uint2(v, __float_as_uint(gain));
}

// One warp per block with candidates: exact dot product and accurate energy of every candidate window, and a
// speculative refine + demodulation + CRC of the windows a walk that enters the block in Searching state would
// detect. All of it is independent of the rest of the capture, so it runs at full occupancy; the per-stream
// walk then only decides (the drift bound depends on where ITS search started) and links the results.
#ifndef TMK_CLUSTER_CTAS
#define TMK_CLUSTER_CTAS 5  // measured at C2, round 1: 4 -> 3.66 ms, 5 -> 3.53, 6 -> 3.50; final kernel: 4 -> 2.447, 5 -> 2.408, 6 -> 2.538 (7+ do not fit: 34 KB smem)
#endif
__global__ void __launch_bounds__(kWarpsPerCta * 32, TMK_CLUSTER_CTAS)
cluster_kernel(const float* __restrict__ samples, const TmkStreamDev* __restrict__ streams,
               const uint32_t* __restrict__ tile_stream, const TmkPattern* __restrict__ patterns,
               const float4* __restrict__ blk_info, const uint32_t* __restrict__ cand_bits, TmkFast f, uint8_t* payload,
               uint64_t payload_cap, TmkCounters* counters) {
    __shared__ TmkCandEntry ents[kWarpsPerCta][64];
    const uint32_t warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t n_items = *f.cblk_count;
    for (;;) {
        // items are pulled, not strided: whatever share of the cluster CTAs is resident beside the walk makes
        // progress on every listed block
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(f.item_next, 1u);
        item = __shfl_sync(kFull, item, 0);
        if (item >= n_items) break;
        const uint32_t gb = f.cblk_list[item];
        const uint32_t sidx = tile_stream[gb / TMK_TILE_BLKS];
        const TmkStreamDev sd = streams[sidx];
        Ctx c;
        init_ctx(c, samples, sd, patterns, blk_info, cand_bits, f, lane, warp);
        const uint64_t bb = uint64_t(gb) - sd.blk_base;
        if (c.P > uint32_t(kFastMaxP)) {  // long preambles keep the walk's own evaluation
            if (lane == 0) publish_blk(f.blk_res + gb, TMK_ENT_INVALID, 0.f);
            continue;
        }
        const uint64_t wb0 = bb * TMK_BLK;
        block_prefix(c, bb);
        const uint32_t myword = lane < TMK_BLK_WORDS ? c.cand_bits[bb * TMK_BLK_WORDS + lane] : 0u;
        uint32_t n_ent = 0;
#pragma unroll 1
        for (uint32_t r = 0; r < TMK_BLK_WORDS; ++r) {
            const uint32_t word = __shfl_sync(kFull, myword, int(r));
            if (word == 0) continue;
            const uint32_t wl = r * 32 + lane;
            const uint64_t w = wb0 + wl;
            const bool valid = ((word >> lane) & 1u) && w < c.nwin;
            TmkCandEntry e;
            e.wl = uint16_t(wl);
            e.status = TMK_ENT_NOSPEC;
            e.consumed = 0;
            e.fs_off = 0;
            e.len = 0;
            e.pad_ = 0;
            e.D = 0.f;
            e.Ec = 0.f;
            e.spec = TMK_ENT_INVALID;
            if (valid) e.Ec = __double2float_rn(c.sc->Q[wl + c.P] - c.sc->Q[wl]);
            const uint32_t m = __ballot_sync(kFull, valid);
            if (c.P > uint32_t(kXsMaxP)) {
                if (valid) e.D = dot_avx_order(c, c.x + w);
            } else
            // exact AVX-order dot products, four candidate windows per round (8 lanes = the 8 AVX lanes of a window)
            for (uint32_t rem = m; rem;) {
                uint32_t t = rem;
                for (int k = 0; k < (lane >> 3); ++k) t &= t - 1;  // this group's candidate: the (lane / 8)-th set bit
                const float Dg = dot_avx_coop(c, r * 32 + (t ? uint32_t(__ffs(t) - 1) : 0u), t != 0);
#pragma unroll
                for (int k = 0; k < 4 && rem; ++k) {
                    const int j = __ffs(rem) - 1;
                    rem &= rem - 1;
                    const float Dk = __shfl_sync(kFull, Dg, 8 * k);
                    if (lane == j) e.D = Dk;
                }
            }
            const uint32_t idx = n_ent + __popc(m & ((1u << lane) - 1u));
            if (valid && idx < 64) ents[warp][idx] = e;
            n_ent += __popc(m);
        }
        __syncwarp();
        if (n_ent > 63) {
            if (lane == 0) publish_blk(f.blk_res + gb, TMK_ENT_INVALID, 0.f);
            continue;
        }
        // gain = max over the block's windows of (sum of the window energies of the block up to w) / S_w,
        // for fast_block's window-by-window form of the drift bound (NaN propagates and disables it)
        // Cheap form first: (sum of all window energies of the block) / (smallest one) bounds every ratio; only a
        // block with a strong contrast (a quiet gap running into a loud frame) needs the window-by-window maximum.
        float gain = 0.f;
        if (c.C != 0) {
            // chunked feeding (chains of at most C + P steps): fast_block bounds the largest window energy of a chain
            // window by window with  mgain = max_w (largest window energy of the block up to w) / S_w
            float carry = 0.f;
#pragma unroll 1
            for (uint32_t r = 0; r < TMK_BLK_WORDS; ++r) {
                const uint32_t wl = r * 32 + lane;
                const bool valid = wb0 + wl < c.nwin;
                const float S = valid ? __double2float_rn(c.sc->Q[wl + c.P] - c.sc->Q[wl]) : 0.f;
                const float pm = max_nan(warp_incl_scan_maxnan(S, lane), carry);
                carry = __shfl_sync(kFull, pm, 31);
                if (valid) gain = max_nan(gain, S > 0.f ? __fdiv_ru(pm * 1.0001f, S) : (S == 0.f ? __int_as_float(0x7f800000) : S));
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) gain = max_nan(gain, __shfl_xor_sync(kFull, gain, o));
        } else {
            // from the scan's summary alone: (windows of the block) x (largest energy) / (smallest energy) bounds every
            // ratio; the scan's energies are within 128 u (neighbourhood maximum) of the exact ones
            const float4 bi = c.blk_info[bb];
            float nbm = bi.y;
            if (bb > 0) nbm = max_nan(nbm, c.blk_info[bb - 1].y);
            if (bb + 1 < c.nblk) nbm = max_nan(nbm, c.blk_info[bb + 1].y);
            const float lo_e = bi.x - kScanErr * nbm;
            const float quick = lo_e > 0.5f * bi.x ? __fdiv_ru(float(TMK_BLK) * (bi.y + kScanErr * nbm) * 1.001f, lo_e) : __int_as_float(0x7f800000);
            if (quick <= 1.0e5f) gain = quick;
        }
        if (c.C == 0 && !(gain > 0.f)) {
            float ssum = 0.f, smin = __int_as_float(0x7f800000);
            bool bad = false;
#pragma unroll 1
            for (uint32_t r = 0; r < TMK_BLK_WORDS; ++r) {
                const uint32_t wl = r * 32 + lane;
                if (wb0 + wl < c.nwin) {
                    const float S = __double2float_rn(c.sc->Q[wl + c.P] - c.sc->Q[wl]);
                    ssum += S;
                    smin = fminf(smin, S);
                    bad = bad || !(S == S);
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                ssum += __shfl_xor_sync(kFull, ssum, o);
                smin = fminf(smin, __shfl_xor_sync(kFull, smin, o));
            }
            bad = __any_sync(kFull, bad);
            const float cheap = smin > 0.f ? __fdiv_ru(ssum * 1.001f, smin) : __int_as_float(0x7f800000);
            if (!bad && cheap <= 1.0e5f) {
                gain = cheap;
            } else {
                float carry = 0.f;
#pragma unroll 1
                for (uint32_t r = 0; r < TMK_BLK_WORDS; ++r) {
                    const uint32_t wl = r * 32 + lane;
                    const bool valid = wb0 + wl < c.nwin;
                    const float S = valid ? __double2float_rn(c.sc->Q[wl + c.P] - c.sc->Q[wl]) : 0.f;
                    float cs = S * 1.0001f;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const float t = __shfl_up_sync(kFull, cs, o);
                        if (lane >= o) cs += t;
                    }
                    cs += carry;
                    carry = __shfl_sync(kFull, cs, 31);
                    if (valid) gain = max_nan(gain, S > 0.f ? __fdiv_ru(cs, S) : (S == 0.f ? __int_as_float(0x7f800000) : S));
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) gain = max_nan(gain, __shfl_xor_sync(kFull, gain, o));
            }
        }
        // speculative decodes, mimicking a walk that meets the entries in order
        uint64_t skip_until = 0;
        for (uint32_t i = 0; i < n_ent; ++i) {
            const TmkCandEntry e = ents[warp][i];
            const uint64_t w = wb0 + e.wl;
            if (w < skip_until) continue;
            const float corr = ex_corr(e.D, e.Ec, c.pat->pe);
            if (!(corr >= c.thr)) continue;
            // the frame behind the preamble starts its trip from HBM now (the refinement and the header decode run first)
            for (uint32_t i = uint32_t(lane) * 32; i < 2304 && w + c.P + i < c.N; i += 1024)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(c.x + w + c.P + i));
            unsigned int sp = 0;
            if (lane == 0) sp = atomicAdd(&f.pool[1], 1u);
            sp = __shfl_sync(kFull, sp, 0);
            if (sp >= f.spec_cap) break;  // pool exhausted: the walk decodes the rest itself
            spec_decode(c, w, corr, sidx, f.spec + sp, payload, payload_cap, counters);
            __syncwarp();
            const tmk_frame_rec* r = f.spec + sp;
            const uint32_t st = __shfl_sync(kFull, lane == 0 ? r->status : 0u, 0);
            const uint32_t cons = __shfl_sync(kFull, lane == 0 ? r->consumed : 0u, 0);
            if (lane == 0) {  // what a walk needs to go on without reading the record
                ents[warp][i].spec = sp;
                ents[warp][i].status = uint16_t(st);
                ents[warp][i].consumed = cons;
                ents[warp][i].fs_off = int32_t(int64_t(r->frame_start) - int64_t(w));
                ents[warp][i].len = r->len;
            }
            __syncwarp();
            if (st == TMK_FRAME_TRUNCATED) break;
            skip_until = w + cons;
        }
        __syncwarp();
        unsigned int base = 0;
        if (lane == 0) base = n_ent ? atomicAdd(&f.pool[0], n_ent) : 0u;
        base = __shfl_sync(kFull, base, 0);
        if (base + n_ent > f.entry_cap || base + n_ent >= (1u << 26)) {
            if (lane == 0) publish_blk(f.blk_res + gb, TMK_ENT_INVALID, 0.f);
            continue;
        }
        for (uint32_t i = lane; i < n_ent; i += 32) f.entries[base + i] = ents[warp][i];
        __syncwarp();
        if (lane == 0) publish_blk(f.blk_res + gb, (base << 6) | n_ent, gain);
    }
}

#ifndef TMK_WALK_CTAS
// measured at C2 (walk kernel, ms), round 1: 2 CTAs/SM 3.72, 3 -> 3.71, 4 -> 3.99, 5 -> 4.21, 6 -> 4.44. The kernel is the sum of
// the streams' serial latencies; beyond 12 warps per SM every warp slows the others down (instruction-cache misses
// of warps that sit in different stages of 136 KB of code, spills below 168 registers).
// End of round 2 (the cluster pass has taken most of the work): 2 -> 2.92, 3 -> 3.05, 4 -> 3.37 at C2; with 2 CTAs per SM
// (up to 255 registers) Manchester 2.55 -> 2.41, C4 2.69 -> 2.58, C3 (16 384 long captures: throughput, not latency) 36.1 -> 37.0.
#define TMK_WALK_CTAS 2
#endif
// CTAS: 2 for batches whose time is the serial latency of their slowest captures, 3 (168 registers) for batches of many long
// captures, where throughput counts (C3 / C5: 12 288 work items and more)
template <bool SEG, int CTAS>  // SEG: the work items are segments of captures (segmented walk); the plain walk carries none of its state
__global__ void __launch_bounds__(kWarpsPerCta * 32, CTAS)
walk_kernel(const float* __restrict__ samples, const TmkStreamDev* __restrict__ streams, uint32_t s0, uint32_t s1,
            unsigned int* __restrict__ next, TmkFast fast,
            const TmkPattern* __restrict__ patterns, const float4* __restrict__ blk_info,
            const uint32_t* __restrict__ cand_bits, tmk_frame_rec* recs, uint4* rec_ord, uint64_t rec_cap,
            uint8_t* payload, uint64_t payload_cap, uint32_t* stream_cnt, TmkCounters* counters,
            TmkLinkState* link_out, TmkSeg seg) {
    const uint32_t warp = threadIdx.x >> 5;
    constexpr bool segm = SEG;  // (tmk_internal.cuh: segmented walk)
    const int lane_id = threadIdx.x & 31;
    Emit em{recs, rec_ord, rec_cap, payload, payload_cap, counters, 0, 0, 0ull, 0ull};
    // persistent warps: streams differ a lot in cost (idle noise vs dense traffic), so every warp
    // pulls the next stream from a counter instead of owning a fixed one
    for (;;) {
    uint32_t sidx = 0;
    if (lane_id == 0) sidx = s0 + atomicAdd(next, 1u);
    sidx = __shfl_sync(kFull, sidx, 0);
    if (sidx >= s1) {
        for (unsigned long long sl = em.slot_cur + lane_id; sl < em.slot_end; sl += 32)  // reserved, never used
            if (sl < rec_cap) rec_ord[sl].y = kNoRecord;
        return;
    }
    uint32_t real = sidx, ord0 = 0;
    uint64_t q1 = ~0ull;  // pass 1 of a segmented walk: pause at the first search entry at or behind this window
    if (segm) {
        real = seg.items[sidx].stream;
        if (seg.pass == 1) {
            q1 = seg.items[sidx].q1;
        } else {  // pass 2: only walkers that paused at the end of their segment go on
            const TmkSegState* ps = seg.st + sidx;
            if (ps->state != 1) continue;
            ord0 = ps->ord;
        }
    }
    const TmkStreamDev sd = streams[real];
    Ctx c;
    init_ctx(c, samples, sd, patterns, blk_info, cand_bits, fast, lane_id, warp);
    em.stream = sidx;
    em.ord = ord0;
    uint32_t n_ok = 0, n_crc = 0, n_hdr = 0, n_nfu = 0;
    uint32_t seg_exit = 0, sync_item = 0, sync_ord = 0;

    const int lc = c.pat->line_code;
    const uint32_t spl = c.pat->spl;
    const uint32_t hs = c.pat->hdr_samples;
    uint64_t q = sd.init_q;                   // buffer_offset (absolute)
    if (segm) q = seg.pass == 1 ? seg.items[sidx].q0 : seg.st[sidx].q;
    uint64_t avail = c.C ? c.C : c.N;         // samples fed so far
    bool resume = sd.init_fs != TMK_NO_FRAME;  // the previous process_samples call ended in Decoding(frame_start)
    TmkLinkState fin;
    fin.off = q;
    fin.fs = TMK_NO_FRAME;
    fin.corr = 0.f;
    fin.pad_ = 0;
    Detection d;
    for (;;) {
        uint64_t fs;
        if (segm) {  // q is a search entry
            if (seg.pass == 1) {
                if (em.ord >= 1 && em.ord <= TMK_SEG_HEADS && lane_id == 0) seg.st[sidx].head[em.ord - 1] = q;
                if (q >= q1) {
                    seg_exit = 1;
                    break;
                }
            } else {
                // has the walker that owns this position been here? (its first search entries, written by pass 1)
                const TmkSegItem it = seg.items[sidx];
                uint64_t k = q / it.seg_len;
                if (k >= it.n_seg) k = it.n_seg - 1;
                const uint32_t owner = it.first_item + uint32_t(k);
                const TmkSegState* os = seg.st + owner;
                const uint32_t nh = os->n_head;
                bool hit = false;
                if (lane_id == 0) hit = seg.items[owner].q0 == q;
                else if (uint32_t(lane_id) <= nh) hit = os->head[lane_id - 1] == q;
                const uint32_t hm = __ballot_sync(kFull, hit);
                if (hm && owner != sidx) {
                    sync_item = owner;
                    sync_ord = uint32_t(__ffs(hm) - 1);
                    seg_exit = 2;
                    break;
                }
            }
        }
        if (resume) {
            resume = false;
            d.c = sd.init_q;
            d.seed = sd.init_q;
            d.avail = avail;
            d.corr = sd.init_corr;
            d.spec = TMK_ENT_INVALID;
            fs = sd.init_fs;
        } else {
        if (!search(c, q, avail, &d)) {
            // search_for_preamble scanned every window of the buffer: buffer_offset += windows (decoder.rs:236)
            fin.off = (c.N >= q && c.N - q >= c.P) ? c.N - c.P + 1 : q;
            break;
        }
        c.n_det++;
        avail = d.avail;
        if (d.spec != TMK_ENT_INVALID && d.c + c.P - c.pat->sync_len - d.seed >= c.pat->margin &&
            (d.avail >= c.N || d.avail >= d.c + c.P + c.pat->margin)) {
            // the cluster pass already refined, demodulated and checked the frame of this detection (its
            // refinement range [expect - margin, expect + margin] is the walk's because the search started at
            // least `margin` samples before `expect` and, with chunked feeding, the samples fed when the window
            // was examined reach `expect + margin + sync_len`, decoder.rs:179-181): link the verdict, payload
            // bytes are in place
            if (d.status == TMK_ENT_NOSPEC) {  // found through the window-by-window path: the summary is in the record
                const tmk_frame_rec r = ld_spec(c.spec + d.spec);
                d.status = r.status;
                d.fs = r.frame_start;
                d.consumed = r.consumed;
                d.len = r.len;
            }
            if (c.C != 0 && d.status != TMK_FRAME_TRUNCATED) {
                // decode_frame waited for the header, then for the whole frame (decoder.rs:248-254, 289-297)
                const bool hdr_only = d.status == TMK_FRAME_HDR_PARSE || d.status == TMK_FRAME_BAD_LEN;
                const uint64_t need = d.fs + (hdr_only ? uint64_t(hs) : tmk_sfb(lc, spl, (TMK_PHY_HEADER_BYTES + uint64_t(d.len)) * 8));
                while (avail < need && avail < c.N) avail = avail + c.C < c.N ? avail + c.C : c.N;
            }
            {   // the verdict is LINKED, not copied: the ordering kernel gathers the record from the speculative pool
                const unsigned long long slot = take_slot(c.lane, em);
                if (c.lane == 0 && slot < rec_cap) rec_ord[slot] = make_uint4(d.spec, em.ord, sidx, __float_as_uint(d.corr));
            }
            em.ord++;
            if (d.status == TMK_FRAME_OK) n_ok++;
            else if (d.status == TMK_FRAME_CRC_FAIL) n_crc++;
            else if (d.status == TMK_FRAME_NOT_FOR_US) n_nfu++;
            else if (d.status != TMK_FRAME_TRUNCATED) n_hdr++;
            if (d.status == TMK_FRAME_TRUNCATED) {
                fin.off = d.c;
                fin.fs = d.fs;
                fin.corr = d.corr;
                break;
            }
            q = d.c + d.consumed;
            continue;
        }
        {   // pull the head of the frame towards L2 while the refinement and the header decode run
            const uint64_t f0 = d.c + c.P;
            for (uint64_t i = uint64_t(lane_id) * 32; i < 2560 && f0 + i < c.N; i += 1024)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(c.x + f0 + i));
        }
        fs = refine(c, d);
        }
        // ---- decode_frame (decoder.rs:242-356); `avail` grows by one chunk per call
        bool truncated = false;
        while (avail < fs + hs) {
            if (avail >= c.N) { truncated = true; break; }
            avail = avail + c.C < c.N ? avail + c.C : c.N;
        }
        if (truncated) {
            emit(c, em, TMK_FRAME_TRUNCATED, d, fs, 0, false, nullptr, 0);
            fin.off = d.c;
            fin.fs = fs;
            fin.corr = d.corr;
            break;
        }
        const uint8_t* hb = c.sc->bytes;
        const uint32_t hbits = line_decode(c, fs, 8 * TMK_PHY_HEADER_BYTES);
        const bool have7 = (hbits + 7) / 8 >= TMK_PHY_HEADER_BYTES;
        uint8_t hdr[TMK_PHY_HEADER_BYTES];
        for (int i = 0; i < TMK_PHY_HEADER_BYTES; ++i) hdr[i] = have7 ? hb[i] : 0;
        __syncwarp();
        const uint32_t type = hdr[3];
        if (!have7 || (type != TMK_FRAME_TYPE_DATA && type != TMK_FRAME_TYPE_ACK)) {  // :266-273
            emit(c, em, TMK_FRAME_HDR_PARSE, d, fs, hs, have7, hdr, 0);
            n_hdr++;
            q = d.c + hs;
            continue;
        }
        const uint32_t len = (uint32_t(hdr[0]) << 8) | hdr[1];
        if ((type == TMK_FRAME_TYPE_DATA && len == 0) || len > c.max_frame_bytes) {  // :277-286
            emit(c, em, TMK_FRAME_BAD_LEN, d, fs, 1, true, hdr, 0);
            n_hdr++;
            q = d.c + 1;
            continue;
        }
        const uint32_t total_bits = (TMK_PHY_HEADER_BYTES + len) * 8;
        const uint64_t total_samples = tmk_sfb(lc, spl, total_bits);
        while (avail < fs + total_samples) {
            if (avail >= c.N) { truncated = true; break; }
            avail = avail + c.C < c.N ? avail + c.C : c.N;
        }
        if (truncated) {
            emit(c, em, TMK_FRAME_TRUNCATED, d, fs, 0, false, nullptr, 0);
            fin.off = d.c;
            fin.fs = fs;
            fin.corr = d.corr;
            break;
        }
        const uint32_t nbits = line_decode(c, fs, total_bits);
        const uint32_t consumed = c.P + uint32_t(tmk_sfb(lc, spl, nbits));  // :306-309
        q = d.c + consumed;
        if (nbits < total_bits) {  // :311-320
            emit(c, em, TMK_FRAME_SHORT_DECODE, d, fs, consumed, true, hdr, 0);
            n_hdr++;
            continue;
        }
        if (c.local_addr >= 0 && hdr[6] != uint32_t(c.local_addr)) {  // :322-329
            emit(c, em, TMK_FRAME_NOT_FOR_US, d, fs, consumed, true, hdr, len);
            n_nfu++;
            continue;
        }
        const uint32_t crc = crc8_warp(c.sc->bytes + TMK_PHY_HEADER_BYTES, len, c.lane);
        if (crc != hdr[2]) {  // frame.rs:155-159
            emit(c, em, TMK_FRAME_CRC_FAIL, d, fs, consumed, true, hdr, len);
            n_crc++;
            continue;
        }
        emit(c, em, TMK_FRAME_OK, d, fs, consumed, true, hdr, len);
        n_ok++;
    }
    if (c.lane == 0) {
        if (link_out) link_out[sidx] = fin;
        stream_cnt[sidx] = em.ord;
        if (segm) {  // the verdict tallies come from the records that survive the resolve step
            TmkSegState* ps = seg.st + sidx;
            ps->q = q;
            ps->ord = em.ord;
            ps->state = seg_exit;
            ps->sync_item = sync_item;
            ps->sync_ord = sync_ord;
            if (seg.pass == 1) ps->n_head = em.ord < TMK_SEG_HEADS ? em.ord : TMK_SEG_HEADS;
            if (seg.pass == 1 && seg.items[sidx].seg == 0) atomicAdd(&counters->stats[TMK_STAT_SAMPLES], (unsigned long long)c.N);
        } else {
        atomicAdd(&counters->stats[TMK_STAT_SAMPLES], (unsigned long long)c.N);
        atomicAdd(&counters->stats[TMK_STAT_DETECTIONS], (unsigned long long)c.n_det);
        atomicAdd(&counters->stats[TMK_STAT_FRAMES_OK], (unsigned long long)n_ok);
        atomicAdd(&counters->stats[TMK_STAT_CRC_FAIL], (unsigned long long)n_crc);
        atomicAdd(&counters->stats[TMK_STAT_HDR_FAIL], (unsigned long long)n_hdr);
        atomicAdd(&counters->stats[TMK_STAT_NOT_FOR_US], (unsigned long long)n_nfu);
        }
        atomicAdd(&counters->stats[TMK_STAT_EXACT_REPLAYS], (unsigned long long)c.n_replays);
        atomicAdd(&counters->stats[TMK_STAT_CANDIDATES], (unsigned long long)c.n_cands);
        for (int i = 0; i < 4; ++i) atomicAdd(&counters->dbg[i], c.dbg[i]);
        atomicAdd(&counters->dbg[5], (unsigned long long)c.n_par_steps);
        atomicAdd(&counters->dbg[6], (unsigned long long)c.n_par_doubt);
    }
    }  // next stream
}

// ---------------------------------------------------------------- ordering of the records
// exclusive scan of the per-stream record counts (single CTA; n_streams is small next to the samples)
__global__ void __launch_bounds__(1024)
scan_counts_kernel(const uint32_t* __restrict__ cnt, uint64_t* __restrict__ base, uint32_t n) {
    __shared__ uint64_t warp_tot[32];
    __shared__ uint64_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t b0 = 0; b0 < n; b0 += 1024) {
        const uint32_t i = b0 + threadIdx.x;
        const uint64_t v = i < n ? cnt[i] : 0;
        uint64_t s = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t t = __shfl_up_sync(kFull, s, o);
            if (lane >= o) s += t;
        }
        if (lane == 31) warp_tot[warp] = s;
        __syncthreads();
        if (warp == 0) {
            uint64_t w = warp_tot[lane], ws = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint64_t t = __shfl_up_sync(kFull, ws, o);
                if (lane >= o) ws += t;
            }
            warp_tot[lane] = ws - w;  // exclusive
        }
        __syncthreads();
        const uint64_t c0 = carry_s;
        if (i < n) base[i] = c0 + warp_tot[warp] + s - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = c0 + warp_tot[31] + s;
        __syncthreads();
    }
    if (threadIdx.x == 0) base[n] = carry_s;
}

__global__ void order_kernel(const tmk_frame_rec* __restrict__ in, const uint4* __restrict__ ord, uint64_t n,
                             const uint64_t* __restrict__ base, const tmk_frame_rec* __restrict__ spec,
                             tmk_frame_rec* __restrict__ out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 ln = ord[i];
    if (ln.y == kNoRecord) return;  // a slot some warp reserved and never used
    tmk_frame_rec r = ln.x != TMK_ENT_INVALID ? spec[ln.x] : in[i];  // linked verdict of the cluster pass | the walk's own
    r.stream = ln.z;
    r.corr = __uint_as_float(ln.w);
    out[base[ln.z] + ln.y] = r;
}

// Small batches (engine.cu: nothing is known on the host yet): exclusive scan of the per-stream counts AND the scatter in one
// CTA, the number of reserved slots read on the device; records beyond `out_cap` are left out (the host sees the exact
// count and takes the general path then)
__global__ void __launch_bounds__(1024)
count_order_small_kernel(const uint32_t* __restrict__ cnt, uint64_t* __restrict__ base, uint32_t n_streams,
                         const tmk_frame_rec* __restrict__ in, const uint4* __restrict__ ord, const TmkCounters* __restrict__ counters,
                         uint64_t slot_cap, const tmk_frame_rec* __restrict__ spec, tmk_frame_rec* __restrict__ out, uint64_t first_cap,
                         tmk_frame_rec* __restrict__ out_rest, uint64_t out_cap, uint64_t* __restrict__ n_exact) {
    __shared__ uint64_t warp_tot[32];
    __shared__ uint64_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t b0 = 0; b0 < n_streams; b0 += 1024) {
        const uint32_t i = b0 + threadIdx.x;
        const uint64_t v = i < n_streams ? cnt[i] : 0;
        uint64_t s = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t t = __shfl_up_sync(kFull, s, o);
            if (lane >= o) s += t;
        }
        if (lane == 31) warp_tot[warp] = s;
        __syncthreads();
        if (warp == 0) {
            uint64_t w = warp_tot[lane], ws = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint64_t t = __shfl_up_sync(kFull, ws, o);
                if (lane >= o) ws += t;
            }
            warp_tot[lane] = ws - w;  // exclusive
        }
        __syncthreads();
        const uint64_t c0 = carry_s;
        if (i < n_streams) base[i] = c0 + warp_tot[warp] + s - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = c0 + warp_tot[31] + s;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        base[n_streams] = carry_s;
        *n_exact = carry_s;
    }
    __syncthreads();  // (the bases written above are read below by other threads of this CTA)
    const uint64_t n = min(uint64_t(counters->n_recs), slot_cap);
    for (uint64_t i = threadIdx.x; i < n; i += 1024) {
        const uint4 ln = ord[i];
        if (ln.y == kNoRecord) continue;
        const uint64_t pos = base[ln.z] + ln.y;
        if (pos >= out_cap) continue;
        tmk_frame_rec r = ln.x != TMK_ENT_INVALID ? spec[ln.x] : in[i];
        r.stream = ln.z;
        r.corr = __uint_as_float(ln.w);
        if (pos < first_cap) out[pos] = r;  // (the first records sit next to the header and the payload: one copy brings a usual poll over)
        else out_rest[pos - first_cap] = r;
    }
}

// ---- segmented walk: which records count
constexpr uint32_t kSegInvalid = 0xffffffffu;
// one thread per capture: from its walker 0 along the meeting points
__global__ void seg_chain_kernel(const TmkSegItem* __restrict__ items, const TmkSegState* __restrict__ st, uint32_t n_items,
                                 uint32_t* __restrict__ first_valid, uint32_t* __restrict__ n_valid) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_items) return;
    const TmkSegItem it = items[i];
    if (it.seg != 0) return;
    for (uint32_t k = 0; k < it.n_seg; ++k) {
        first_valid[i + k] = kSegInvalid;
        n_valid[i + k] = 0;
    }
    uint32_t cur = i, start = 0;
    for (uint32_t guard = 0; guard <= it.n_seg; ++guard) {
        const TmkSegState s = st[cur];
        first_valid[cur] = start;
        n_valid[cur] = s.ord >= start ? s.ord - start : 0;
        if (s.state != 2 || s.sync_item <= cur || s.sync_item >= i + it.n_seg) break;
        start = s.sync_ord;
        cur = s.sync_item;
    }
}
// verdict tallies of the records that count (the walkers of a segmented walk do not tally: part of their records is dropped)
__global__ void seg_tally_kernel(const tmk_frame_rec* __restrict__ in, const uint4* __restrict__ ord, uint64_t rec_cap,
                                 const tmk_frame_rec* __restrict__ spec, const uint32_t* __restrict__ first_valid,
                                 TmkCounters* counters) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t n = counters->n_recs;  // slots the walkers reserved (the host does not know the number yet)
    uint32_t status = 0xffu;
    if (i < n && i < rec_cap) {
        const uint4 ln = ord[i];
        if (ln.y != kNoRecord) {
            const uint32_t f = first_valid[ln.z];
            if (f != kSegInvalid && ln.y >= f) status = ln.x != TMK_ENT_INVALID ? spec[ln.x].status : in[i].status;
        }
    }
    const uint32_t any = __ballot_sync(kFull, status != 0xffu);
    if (!any) return;
    const uint32_t ok = __ballot_sync(kFull, status == TMK_FRAME_OK), crc = __ballot_sync(kFull, status == TMK_FRAME_CRC_FAIL);
    const uint32_t nfu = __ballot_sync(kFull, status == TMK_FRAME_NOT_FOR_US), tr = __ballot_sync(kFull, status == TMK_FRAME_TRUNCATED);
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&counters->stats[TMK_STAT_DETECTIONS], (unsigned long long)__popc(any));
        if (ok) atomicAdd(&counters->stats[TMK_STAT_FRAMES_OK], (unsigned long long)__popc(ok));
        if (crc) atomicAdd(&counters->stats[TMK_STAT_CRC_FAIL], (unsigned long long)__popc(crc));
        if (nfu) atomicAdd(&counters->stats[TMK_STAT_NOT_FOR_US], (unsigned long long)__popc(nfu));
        const uint32_t hdr = any & ~(ok | crc | nfu | tr);
        if (hdr) atomicAdd(&counters->stats[TMK_STAT_HDR_FAIL], (unsigned long long)__popc(hdr));
    }
}
__global__ void seg_order_kernel(const tmk_frame_rec* __restrict__ in, const uint4* __restrict__ ord, uint64_t n,
                                 const uint64_t* __restrict__ base, const tmk_frame_rec* __restrict__ spec,
                                 const TmkSegItem* __restrict__ items, const uint32_t* __restrict__ first_valid,
                                 tmk_frame_rec* __restrict__ out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 ln = ord[i];
    if (ln.y == kNoRecord) return;
    const uint32_t f = first_valid[ln.z];
    if (f == kSegInvalid || ln.y < f) return;  // before the walker's meeting point, or a walker that was passed over
    tmk_frame_rec r = ln.x != TMK_ENT_INVALID ? spec[ln.x] : in[i];
    r.stream = items[ln.z].stream;
    r.corr = __uint_as_float(ln.w);
    out[base[ln.z] + (ln.y - f)] = r;
}

}  // namespace

// parity probe of the exact replay (tmk_replay_probe): E_ref(w[i]) of a search seeded at `seed`, through the very
// function the walk uses (cached chain position, lane-parallel rounds, literal rounds)
__global__ void __launch_bounds__(32)
replay_probe_kernel(const float* __restrict__ samples, const TmkStreamDev* __restrict__ streams,
                    const TmkPattern* __restrict__ patterns, uint64_t seed, const uint64_t* __restrict__ ws, uint32_t n,
                    float* __restrict__ out, unsigned long long* __restrict__ counts) {
    Ctx c;
    TmkFast nofast;
    memset(&nofast, 0, sizeof nofast);
    init_ctx(c, samples, streams[0], patterns, nullptr, nullptr, nofast, int(threadIdx.x), 0);
    for (uint32_t i = 0; i < n; ++i) {
        const float e = exact_energy(c, seed, ws[i]);
        if (threadIdx.x == 0) out[i] = e;
    }
    if (threadIdx.x == 0) {
        counts[0] = c.dbg[2];        // literal steps
        counts[1] = c.n_par_steps;   // steps covered by lane-parallel rounds
        counts[2] = c.n_par_doubt;   // segments rolled literally inside them
    }
}

int tmk_walk_init() {
    uint8_t tab[kMaxFrameBytesCap + 8];
    uint32_t v = 1;  // x^0
    for (int k = 0; k < kMaxFrameBytesCap + 8; ++k) {
        tab[k] = uint8_t(v);
        for (int b = 0; b < 8; ++b) v = (v & 0x80u) ? ((v << 1) ^ 0x07u) & 0xffu : (v << 1) & 0xffu;  // times x^8
    }
    return cudaMemcpyToSymbol(g_crc_pow, tab, sizeof tab) == cudaSuccess ? 0 : -1;
}

int tmk_walk_max_frame_bytes() { return kMaxFrameBytesCap; }

// base[i] = sum of cnt[0..i), base[n] = total
int tmk_launch_excl_scan(const uint32_t* cnt, uint64_t* base, uint32_t n, cudaStream_t st) {
    scan_counts_kernel<<<1, 1024, 0, st>>>(cnt, base, n);
    return 1;
}

int tmk_launch_cluster(const float* samples, const TmkStreamDev* streams, const uint32_t* tile_stream,
                       const TmkPattern* patterns, const float4* blk_info, const uint32_t* cand_bits, TmkFast f,
                       uint8_t* payload, uint64_t payload_cap, TmkCounters* counters, bool beside_walk, cudaStream_t st) {
    // beside the walk: half of the SMs' CTA slots: the walk (launched right after, on another stream) takes the other half at
    // once and the rest as these CTAs retire
    static int max_ctas = 0;
    if (max_ctas == 0) {
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cluster_kernel, kWarpsPerCta * 32, 0);
        max_ctas = sms * (per_sm > 0 ? per_sm : 1);
    }
    int grid = max_ctas;
    if (beside_walk) {
        int share = 2;
        if (const char* ev = getenv("TMK_CLUSTER_CTAS")) share = std::max(1, atoi(ev));
        grid = std::max(1, max_ctas / 4 * share);
    }
    cluster_kernel<<<grid, kWarpsPerCta * 32, 0, st>>>(samples, streams, tile_stream, patterns, blk_info, cand_bits, f,
                                                           payload, payload_cap, counters);
    return 1;
}

int tmk_launch_walk(const float* samples, const TmkStreamDev* streams, uint32_t s0, uint32_t s1, unsigned int* next,
                    TmkFast fast,
                    const TmkPattern* patterns, const float4* blk_info, const uint32_t* cand_bits,
                    tmk_frame_rec* recs, uint4* rec_ord, uint64_t rec_cap, uint8_t* payload,
                    uint64_t payload_cap, uint32_t* stream_cnt, TmkCounters* counters, TmkLinkState* link_out,
                    cudaStream_t st, TmkSeg seg) {
    if (s1 <= s0) return 0;
    const uint32_t n_streams = s1 - s0;
    static int max_ctas[2] = {0, 0};  // [0]: TMK_WALK_CTAS per SM, [1]: 3 per SM
    if (max_ctas[0] == 0) {
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, walk_kernel<false, TMK_WALK_CTAS>, kWarpsPerCta * 32, 0);
        max_ctas[0] = sms * (per_sm > 0 ? per_sm : 1);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, walk_kernel<false, 3>, kWarpsPerCta * 32, 0);
        max_ctas[1] = sms * (per_sm > 0 ? per_sm : 1);
    }
    const bool many = seg.items == nullptr && n_streams >= 12288;
    uint32_t grid = (n_streams + kWarpsPerCta - 1) / kWarpsPerCta;
    if (grid > uint32_t(max_ctas[many ? 1 : 0])) grid = uint32_t(max_ctas[many ? 1 : 0]);
    if (seg.items != nullptr)
        walk_kernel<true, TMK_WALK_CTAS><<<grid, kWarpsPerCta * 32, 0, st>>>(samples, streams, s0, s1, next, fast, patterns, blk_info, cand_bits,
                                                                             recs, rec_ord, rec_cap, payload, payload_cap, stream_cnt, counters,
                                                                             link_out, seg);
    else if (many)
        walk_kernel<false, 3><<<grid, kWarpsPerCta * 32, 0, st>>>(samples, streams, s0, s1, next, fast, patterns, blk_info, cand_bits, recs,
                                                                  rec_ord, rec_cap, payload, payload_cap, stream_cnt, counters, link_out, seg);
    else
        walk_kernel<false, TMK_WALK_CTAS><<<grid, kWarpsPerCta * 32, 0, st>>>(samples, streams, s0, s1, next, fast, patterns, blk_info, cand_bits,
                                                                              recs, rec_ord, rec_cap, payload, payload_cap, stream_cnt, counters,
                                                                              link_out, seg);
    return 1;
}

int tmk_launch_seg_resolve(const TmkSegItem* items, const TmkSegState* st, uint32_t n_items, uint32_t* first_valid,
                           uint32_t* n_valid, const tmk_frame_rec* recs, const uint4* rec_ord, uint64_t n_slots, uint64_t rec_cap,
                           const tmk_frame_rec* spec, TmkCounters* counters, cudaStream_t stream) {
    seg_chain_kernel<<<(n_items + 127) / 128, 128, 0, stream>>>(items, st, n_items, first_valid, n_valid);
    if (n_slots == 0) return 1;
    seg_tally_kernel<<<uint32_t((n_slots + 255) / 256), 256, 0, stream>>>(recs, rec_ord, rec_cap, spec, first_valid, counters);
    return 2;
}

// the chain of meeting points alone (the chirp receiver's segmented walk, ask.cu, tallies nothing)
int tmk_launch_seg_chain(const TmkSegItem* items, const TmkSegState* st, uint32_t n_items, uint32_t* first_valid, uint32_t* n_valid,
                         cudaStream_t stream) {
    seg_chain_kernel<<<(n_items + 127) / 128, 128, 0, stream>>>(items, st, n_items, first_valid, n_valid);
    return 1;
}

int tmk_launch_seg_order(const tmk_frame_rec* recs_in, const uint4* rec_ord, uint64_t n_slots, const uint64_t* base,
                         const tmk_frame_rec* spec, const TmkSegItem* items, const uint32_t* first_valid,
                         tmk_frame_rec* recs_out, cudaStream_t stream) {
    if (n_slots == 0) return 0;
    seg_order_kernel<<<uint32_t((n_slots + 255) / 256), 256, 0, stream>>>(recs_in, rec_ord, n_slots, base, spec, items, first_valid,
                                                                        recs_out);
    return 1;
}

int tmk_launch_replay_probe(const float* samples, const TmkStreamDev* streams, const TmkPattern* patterns, uint64_t seed,
                            const uint64_t* ws, uint32_t n, float* out, unsigned long long* counts, cudaStream_t st) {
    replay_probe_kernel<<<1, 32, 0, st>>>(samples, streams, patterns, seed, ws, n, out, counts);
    return 1;
}

int tmk_launch_order_small(const uint32_t* stream_cnt, uint64_t* stream_base, uint32_t n_streams, const tmk_frame_rec* recs_in,
                           const uint4* rec_ord, const TmkCounters* counters, uint64_t slot_cap, const tmk_frame_rec* spec,
                           tmk_frame_rec* recs_out, uint64_t first_cap, tmk_frame_rec* recs_out_rest, uint64_t out_cap, uint64_t* n_exact,
                           cudaStream_t st) {
    count_order_small_kernel<<<1, 1024, 0, st>>>(stream_cnt, stream_base, n_streams, recs_in, rec_ord, counters, slot_cap, spec, recs_out,
                                                 first_cap, recs_out_rest, out_cap, n_exact);
    return 1;
}

int tmk_launch_order(const tmk_frame_rec* recs_in, const uint4* rec_ord, uint64_t n_recs,
                     const uint32_t* stream_cnt, uint64_t* stream_base, uint32_t n_streams,
                     const tmk_frame_rec* spec, tmk_frame_rec* recs_out, cudaStream_t st) {
    tmk_launch_excl_scan(stream_cnt, stream_base, n_streams, st);
    if (n_recs == 0) return 1;
    const uint32_t grid = uint32_t((n_recs + 255) / 256);
    order_kernel<<<grid, 256, 0, st>>>(recs_in, rec_ord, n_recs, stream_base, spec, recs_out);
    return 2;
}
