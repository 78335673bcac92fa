// scan_fast.cu — preamble scan specialised for the reference's shipped constants
// (SAMPLES_PER_LEVEL = 3, PREAMBLE_PATTERN_BYTES = 2; src/utils/consts.rs:30-34), the hot kernel
// of the engine: one pass over HBM, 4 B per sample, everything else stays on chip.
//
// It replaces the per-sample loop of PhyDecoder::search_for_preamble (src/phy/decoder.rs:138-147,
// P multiply-adds + sqrt + divide per sample) as a FILTER (see scan_generic.cu for the contract):
// candidate bit per window, per-block (min E, max E, max corr^2, #candidates).
//
// Design (B200):
//  * one WARP owns a tile of 3840 windows (8 blocks of 480); a lane owns a strip of 120 windows.
//  * the lane's 120 + P samples are staged in shared memory by ONE bulk async copy per lane
//    (cp.async.bulk -> UBLKCP, completion on an mbarrier): no staging instructions, no __syncthreads,
//    rows pitched at an odd multiple of 16 B so the lane-strided LDS.128 are conflict free.
//  * the strip is streamed through registers: level sums y[m] = x[m]+x[m+1]+x[m+2] are formed once
//    per sample and folded into per-window accumulators that live in a register ring, using the
//    run structure of the +-1 preamble:
//       4B5B        --++-++--++--+-++-++          11 adds per window instead of 60 multiply-adds
//       Manchester  (+-)(+-)(-+)(-+)... factored   8 adds per window instead of 96
//    the window energy rolls (one FFMA in, one FFMA out).
//  * nothing here is a dense contraction: no tensor cores, fp32 pipe + issue slots are the bound
//    next to HBM (DESIGN.md has the instruction budget).
#ifdef TMK_SCAN_ARM_A  // the default build runs the tensor-core scan (scan_tc.cu); this kernel is the A arm of the A/B
#include "tmk_internal.cuh"

namespace {

constexpr int kS = 120;      // windows per lane strip
constexpr int kMaxWarps = 4; // warps per CTA (each works on its own tiles): Shape<LC>::kWarps <= kMaxWarps
constexpr unsigned kFull = 0xffffffffu;

template <int LC> struct Shape;
template <> struct Shape<TMK_LINE_4B5B> {
    static constexpr int L = 20, P = 60;
    static constexpr int kCopy = 180;  // floats copied per lane row (>= kS + P - 1, multiple of 4)
    static constexpr int kRow = 180;   // row pitch in floats: 720 B = 45 x 16 B (odd: conflict-free LDS.128)
    static constexpr int kWarps = 4;   // 92 KB of rows per CTA, 2 CTAs per SM
};
template <> struct Shape<TMK_LINE_MANCHESTER> {
    static constexpr int L = 32, P = 96;
    static constexpr int kCopy = 216;
    static constexpr int kRow = 220;   // 880 B = 55 x 16 B
    static constexpr int kWarps = 4;   // 113 KB of rows per CTA, 2 CTAs per SM
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1-D bulk async copy global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// asynchronous HBM -> L2 prefetch of a byte range (no completion to wait for)
__device__ __forceinline__ void l2_prefetch(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

struct LaneOut {
    float emin, emax, rmax;
    uint32_t ncand;
};

// static ring index: (i mod R) for i that may be negative
template <int R> __device__ __forceinline__ constexpr int rix(int i) { return ((i % R) + R) % R; }

// finish a window of the strip (branch free): block statistics and the candidate predicate.
// Only tail tiles (FULL == false) have windows to mask: past the end of the stream.
template <bool FULL>
__device__ __forceinline__ bool finish_window(float D, float E, float kap, bool valid, LaneOut& o) {
    const float dp = fmaxf(D, 0.0f);
    if (FULL) {
        const float r = dp * dp * rcp_approx(E);  // P * corr^2 (NaN for silence: never a candidate)
        o.emin = fminf(o.emin, E);
        o.emax = max_nan(o.emax, E);              // a NaN sample stays visible in max E
        o.rmax = fmaxf(o.rmax, r);
        return r > kap;
    }
    const float Ev = valid ? E : __int_as_float(0x7fc00000);  // NaN: ignored by fminf / fmaxf, never a candidate
    const float r = dp * dp * rcp_approx(Ev);
    o.emin = fminf(o.emin, Ev);
    o.emax = max_nan(o.emax, valid ? E : 0.0f);
    o.rmax = fmaxf(o.rmax, r);
    return r > kap;
}

// candidate bits of one loop iteration (U steps): rare, so the bits are flushed with atomics
template <int U, int LAT>
__device__ __forceinline__ void flush_cands(uint64_t cm0, uint64_t cm1, int k, int nvalid, uint32_t* __restrict__ bits,
                                            uint32_t bit0, LaneOut& o) {
    // bit j of (cm1:cm0) is step j of iteration k, i.e. window n = U * k + j - LAT
#pragma unroll 1
    for (int h = 0; h < 2; ++h) {
        uint64_t cm = h ? cm1 : cm0;
        while (cm) {
            const int j = __ffsll((long long)cm) - 1 + 64 * h;
            cm &= cm - 1;
            const int n = U * k + j - LAT;
            if (n >= 0 && n < nvalid) {
                const uint32_t b = bit0 + uint32_t(n);
                atomicOr(bits + (b >> 5), 1u << (b & 31));
                o.ncand++;
            }
        }
    }
}

// The strips run in two phases. Phase A (the first P - 3 steps) only warms the rings up: no window is
// complete yet, so it carries no finish code. Phase B finishes one window per step, all of them inside
// the strip, so nothing has to be masked (tail tiles mask by `n < nvalid`). For 4B5B phase B is a ROLLED
// loop of 2 x 60 steps (every ring size divides 60, so all register indices stay static); the code that
// every warp executes is ~2.6 K instructions and stays resident in the instruction cache (the first
// version, 177 fully unrolled steps with a branch per window, was 148 KB of straight-line code).

// ---------------------------------------------------------------- 4B5B: levels --++-++--++--+-++-++
// D(n) = -d[n] - y[n+12] + p2[n+15] - d[n+21] - p2[n+33] + y[n+39] - y[n+42] + p2[n+45] - y[n+51] + p2[n+54]
//   y[m] = x[m]+x[m+1]+x[m+2]   p2[m] = y[m]+y[m+3]   d[m] = p2[m]-p2[m+6]        (indices in samples)
template <bool FULL>
__device__ __forceinline__ void strip_4b5b(const float* __restrict__ row, float kap, int nvalid,
                                           uint32_t* __restrict__ bits, uint32_t bit0, LaneOut& o) {
    constexpr int U = 60, LAT = 57;  // steps per phase-B iteration; steps from a window's first sample to its finish
    static_assert(2 * U == kS && U % 12 == 0 && U <= 64, "strip steps");
    float acc[60], yb[4], p2b[12], db[4], xr[12], xo[4];
#pragma unroll
    for (int i = 0; i < 12; ++i) p2b[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) yb[i] = db[i] = 0.f;
    float e = 0.0f;
    {
        const float4 v = *reinterpret_cast<const float4*>(row);
        xr[0] = v.x; xr[1] = v.y; xr[2] = v.z; xr[3] = v.w;
        e = fmaf(v.x, v.x, e);
        e = fmaf(v.y, v.y, e);
    }
    // one step of the recurrences at sample m = base + mm (mm static); `rb` = row + base
#define TMK_STEP_4B5B(rb, mm, PHASE_B)                                                                      \
    if ((((mm) + 2) & 3) == 0) {                                                                            \
        const float4 v = *reinterpret_cast<const float4*>((rb) + (mm) + 2);                                 \
        xr[rix<12>((mm) + 2)] = v.x; xr[rix<12>((mm) + 3)] = v.y;                                           \
        xr[rix<12>((mm) + 4)] = v.z; xr[rix<12>((mm) + 5)] = v.w;                                           \
    }                                                                                                       \
    const float x2 = xr[rix<12>((mm) + 2)];                                                                 \
    const float y = (xr[rix<12>(mm)] + xr[rix<12>((mm) + 1)]) + x2;                                         \
    e = fmaf(x2, x2, e);                                                                                    \
    const float p2 = yb[rix<4>((mm) - 3)] + y;                                                              \
    yb[rix<4>(mm)] = y;                                                                                     \
    const float d = p2b[rix<12>((mm) - 9)] - p2;                                                            \
    p2b[rix<12>((mm) - 3)] = p2;                                                                            \
    if (PHASE_B || (mm) >= 12) acc[rix<60>((mm) - 12)] = -db[rix<4>((mm) - 12)] - y;                        \
    db[rix<4>((mm) - 9)] = d;                                                                               \
    if (PHASE_B || (mm) >= 18) acc[rix<60>((mm) - 18)] += p2;                                               \
    if (PHASE_B || (mm) >= 30) acc[rix<60>((mm) - 30)] -= d;                                                \
    if (PHASE_B || (mm) >= 36) acc[rix<60>((mm) - 36)] -= p2;                                               \
    if (PHASE_B || (mm) >= 39) acc[rix<60>((mm) - 39)] += y;                                                \
    if (PHASE_B || (mm) >= 42) acc[rix<60>((mm) - 42)] -= y;                                                \
    if (PHASE_B || (mm) >= 48) acc[rix<60>((mm) - 48)] += p2;                                               \
    if (PHASE_B || (mm) >= 51) acc[rix<60>((mm) - 51)] -= y;

    // ---- phase A: samples 0 .. 56
#pragma unroll
    for (int m = 0; m < LAT; ++m) {
        TMK_STEP_4B5B(row, m, false)
        (void)p2;
    }
    // ---- phase B: samples 57 .. 176, window n = m - 57 finishes at every step
#pragma unroll 1
    for (int k = 0; k < 2; ++k) {
        const float* __restrict__ rk = row + U * k;
        uint64_t cm = 0;
#pragma unroll
        for (int j = 0; j < U; ++j) {
            TMK_STEP_4B5B(rk, LAT + j, true)
            const float D = acc[rix<60>(j)] + p2;  // window n = U * k + j: x[n .. n+59] have entered e
            const bool valid = FULL || (U * k + j) < nvalid;
            if (finish_window<FULL>(D, e, kap, valid, o)) cm |= 1ull << j;
            if ((j & 3) == 0) {
                const float4 v = *reinterpret_cast<const float4*>(rk + j);
                xo[0] = v.x; xo[1] = v.y; xo[2] = v.z; xo[3] = v.w;
            }
            const float xl = xo[j & 3];
            e = fmaf(-xl, xl, e);  // x[n] leaves
        }
        if (cm) flush_cands<U, 0>(cm, 0, k, FULL ? kS : nvalid, bits, bit0, o);
    }
#undef TMK_STEP_4B5B
}

// ---------------------------------------------------------------- Manchester, bits 0011 0011 0101 1010
// D(n) = v[n] + v[n+24] + t[n+48] - t[n+72]
//   z[m] = y[m]-y[m+3]  w[m] = z[m]+z[m+6]  u[m] = z[m]-z[m+6]  v[m] = w[m]-w[m+12]  t[m] = u[m]+u[m+12]
// (phase B is 120 straight-line steps: the accumulator ring needs 73 slots, more than any divisor of 120
//  below 120 offers)
template <bool FULL>
__device__ __forceinline__ void strip_manchester(const float* __restrict__ row, float kap, int nvalid,
                                                 uint32_t* __restrict__ bits, uint32_t bit0, LaneOut& o) {
    constexpr int LAT = 93;
    float acc[80], yb[4], zb[8], wb[16], ub[16], xr[8], xo[4];
#pragma unroll
    for (int i = 0; i < 16; ++i) wb[i] = ub[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) zb[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) yb[i] = 0.f;
    float e = 0.0f;
    {
        const float4 v = *reinterpret_cast<const float4*>(row);
        xr[0] = v.x; xr[1] = v.y; xr[2] = v.z; xr[3] = v.w;
        e = fmaf(v.x, v.x, e);
        e = fmaf(v.y, v.y, e);
    }
    uint64_t cm0 = 0, cm1 = 0;
#pragma unroll
    for (int m = 0; m < kS + LAT; ++m) {
        if (((m + 2) & 3) == 0) {
            const float4 v = *reinterpret_cast<const float4*>(row + m + 2);
            xr[(m + 2) & 7] = v.x; xr[(m + 3) & 7] = v.y; xr[(m + 4) & 7] = v.z; xr[(m + 5) & 7] = v.w;
        }
        const float x2 = xr[(m + 2) & 7];
        const float y = (xr[m & 7] + xr[(m + 1) & 7]) + x2;
        e = fmaf(x2, x2, e);
        const float z = yb[(m + 1) & 3] - y;  // z[m-3]
        yb[m & 3] = y;
        const float zo = zb[(m + 7) & 7];     // z[m-9]
        const float w = zo + z, u = zo - z;   // w[m-9], u[m-9]
        zb[(m + 5) & 7] = z;
        const float vv = wb[(m + 11) & 15] - w;  // v[m-21], t[m-21]
        const float tt = ub[(m + 11) & 15] + u;
        wb[(m + 7) & 15] = w;
        ub[(m + 7) & 15] = u;
        if (m >= 21) acc[(m - 21) % 80] = vv;
        if (m >= 45) acc[(m - 45) % 80] += vv;
        if (m >= 69) acc[(m - 69) % 80] += tt;
        if (m >= LAT) {
            const int n = m - LAT;  // window n is complete
            const float D = acc[n % 80] - tt;
            const bool valid = FULL || n < nvalid;
            if (finish_window<FULL>(D, e, kap, valid, o)) {
                if (n < 64) cm0 |= 1ull << (n & 63);
                else cm1 |= 1ull << (n & 63);
            }
            if ((n & 3) == 0) {
                const float4 v = *reinterpret_cast<const float4*>(row + n);
                xo[0] = v.x; xo[1] = v.y; xo[2] = v.z; xo[3] = v.w;
            }
            const float xl = xo[n & 3];
            e = fmaf(-xl, xl, e);
        }
    }
    if (cm0 | cm1) flush_cands<128, 0>(cm0, cm1, 0, FULL ? kS : nvalid, bits, bit0, o);
}

template <int LC>
__global__ void __launch_bounds__(Shape<LC>::kWarps * 32, 2)
scan_fast_kernel(const float* __restrict__ samples, uint64_t total_samples, const TmkStreamDev* __restrict__ streams,
                 const uint32_t* __restrict__ tile_stream, uint32_t tile0, uint32_t n_tiles,
                 const TmkPattern* __restrict__ patterns, float4* __restrict__ blk_info, uint32_t* __restrict__ cand_bits,
                 uint32_t* __restrict__ cblk_list, unsigned int* __restrict__ cblk_count) {
    using Sh = Shape<LC>;
    extern __shared__ __align__(128) float smem[];
    __shared__ __align__(8) uint64_t mbar[kMaxWarps];
    constexpr int kWarps = Sh::kWarps;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* row = smem + size_t(warp * 32 + lane) * Sh::kRow;
    uint64_t* bar = &mbar[warp];
    if (lane == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    uint32_t parity = 0;
    for (uint32_t tl = blockIdx.x * kWarps + warp; tl < n_tiles; tl += gridDim.x * kWarps) {
        const uint32_t tile = tile0 + tl;  // tiles [tile0, tile0 + n_tiles) of the batch
        const uint32_t sidx = tile_stream[tile];
        const TmkStreamDev sd = streams[sidx];
        const TmkPattern* pat = patterns + sd.pattern;
        if (pat->line_code != LC || pat->spl != 3 || pat->preamble_bytes != 2) continue;  // another kernel's stream
        const uint64_t n = sd.n_samples;
        const uint64_t nwin = n - Sh::P + 1;
        const uint64_t w0 = uint64_t(tile - sd.tile_base) * TMK_TILE;  // first window of the tile
        const uint64_t ls = w0 + uint64_t(lane) * kS;                  // first window / sample of the lane strip
        const uint64_t abs0 = sd.offset + ls;                          // ... in the sample buffer
        // ---- stage the lane rows: one bulk copy per lane whose row lies inside the buffer
        const bool bulk = abs0 + Sh::kCopy <= total_samples;
        const uint32_t nbulk = __popc(__ballot_sync(kFull, bulk));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic reads of the rows are done
        if (lane == 0 && nbulk) mbar_expect_tx(bar, nbulk * Sh::kCopy * 4u);
        __syncwarp();
        if (bulk) {
            bulk_g2s(row, samples + abs0, Sh::kCopy * 4u, bar);
        } else {
            for (int i = 0; i < Sh::kCopy; ++i) row[i] = abs0 + i < total_samples ? samples[abs0 + i] : 0.0f;
        }
        if (nbulk) {
            mbar_wait(bar, parity);
            parity ^= 1u;
        }
        __syncwarp();
#ifndef TMK_SCAN_NO_PREFETCH
        {   // the rows of this warp's NEXT tile start their trip from HBM to L2 now, under the strip below
            const uint32_t tn = tl + gridDim.x * kWarps;
            if (tn < n_tiles) {
                const uint32_t tile_n = tile0 + tn, sn = tile_stream[tile_n];
                const uint64_t an = streams[sn].offset + uint64_t(tile_n - streams[sn].tile_base) * TMK_TILE;
                constexpr uint32_t kSpan = TMK_TILE + Sh::kCopy - kS;  // the 32 rows are one contiguous range
                if (lane == 0 && an + kSpan <= total_samples) l2_prefetch(samples + an, kSpan * 4u);
            }
        }
#endif
        // ---- stream the strip
        const float thr_c = sd.threshold * TMK_CAND_RATIO;
        const float kap = thr_c * thr_c * float(Sh::P);
        const uint64_t blk0 = sd.blk_base + w0 / TMK_BLK;  // first block of the tile
        uint32_t* bits = cand_bits + blk0 * TMK_BLK_WORDS;
        const uint32_t bit0 = uint32_t(lane) * kS;
        LaneOut o{__int_as_float(0x7f800000), 0.0f, 0.0f, 0u};
        const bool full = w0 + TMK_TILE <= nwin;
        int nvalid = kS;
        if (!full) nvalid = ls >= nwin ? 0 : (nwin - ls < uint64_t(kS) ? int(nwin - ls) : kS);
        if (LC == TMK_LINE_4B5B) {
            if (full) strip_4b5b<true>(row, kap, nvalid, bits, bit0, o);
            else strip_4b5b<false>(row, kap, nvalid, bits, bit0, o);
        } else {
            if (full) strip_manchester<true>(row, kap, nvalid, bits, bit0, o);
            else strip_manchester<false>(row, kap, nvalid, bits, bit0, o);
        }
        // ---- block statistics: 4 lanes = one block of 480 windows
        float ncf = float(o.ncand);
#pragma unroll
        for (int s = 1; s <= 2; s <<= 1) {
            o.emin = fminf(o.emin, __shfl_xor_sync(kFull, o.emin, s));
            o.emax = max_nan(o.emax, __shfl_xor_sync(kFull, o.emax, s));
            o.rmax = fmaxf(o.rmax, __shfl_xor_sync(kFull, o.rmax, s));
            ncf += __shfl_xor_sync(kFull, ncf, s);
        }
        if ((lane & 3) == 0 && w0 + uint64_t(lane >> 2) * TMK_BLK < nwin) {
            blk_info[blk0 + (lane >> 2)] = make_float4(o.emin, o.emax, o.rmax * (1.0f / float(Sh::P)), ncf);
            if (ncf > 0.f && cblk_list) cblk_list[atomicAdd(cblk_count, 1u)] = uint32_t(blk0 + (lane >> 2));
        }
        __syncwarp();
    }
}

template <int LC>
void launch_fast(const float* samples, uint64_t total_samples, const TmkStreamDev* streams, const uint32_t* tile_stream,
                 uint32_t tile0, uint32_t n_tiles, const TmkPattern* patterns, float4* blk_info, uint32_t* cand_bits,
                 uint32_t* cblk_list, unsigned int* cblk_count, cudaStream_t st) {
    static int grid_cap = 0;
    constexpr int kWarps = Shape<LC>::kWarps;
    const size_t smem = size_t(kWarps) * 32 * Shape<LC>::kRow * sizeof(float);
    if (grid_cap == 0) {
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaFuncSetAttribute(scan_fast_kernel<LC>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, scan_fast_kernel<LC>, kWarps * 32, smem);
        grid_cap = sms * (per_sm > 0 ? per_sm : 1);
    }
    uint32_t grid = (n_tiles + kWarps - 1) / kWarps;
    if (grid > uint32_t(grid_cap)) grid = uint32_t(grid_cap);
    scan_fast_kernel<LC><<<grid, kWarps * 32, smem, st>>>(samples, total_samples, streams, tile_stream, tile0, n_tiles,
                                                        patterns, blk_info, cand_bits, cblk_list, cblk_count);
}

}  // namespace

// kinds: bit 0 = some stream is 4B5B/3/2, bit 1 = Manchester/3/2, bit 2 = any other shape (generic kernel)
int tmk_launch_scan(const float* samples, uint64_t total_samples, const TmkStreamDev* streams, uint32_t n_streams,
                    const uint32_t* tile_stream, uint32_t tile0, uint32_t n_tiles, const TmkPattern* patterns,
                    uint32_t kinds, float4* blk_info, uint32_t* cand_bits, uint32_t* cblk_list, unsigned int* cblk_count,
                    cudaStream_t st, void* tile_desc, float* dbg_dot, bool build_dir) {
    (void)dbg_dot;
    (void)build_dir;
    (void)tile_desc;
    (void)n_streams;
    if (n_tiles == 0) return 0;
    int launches = 0;
    if (kinds & 1u) {
        launch_fast<TMK_LINE_4B5B>(samples, total_samples, streams, tile_stream, tile0, n_tiles, patterns, blk_info, cand_bits,
                                   cblk_list, cblk_count, st);
        ++launches;
    }
    if (kinds & 2u) {
        launch_fast<TMK_LINE_MANCHESTER>(samples, total_samples, streams, tile_stream, tile0, n_tiles, patterns, blk_info,
                                         cand_bits, cblk_list, cblk_count, st);
        ++launches;
    }
    if (kinds & 4u) launches += tmk_launch_scan_generic(samples, streams, tile_stream, tile0, n_tiles, patterns, blk_info, cand_bits,
                                                            cblk_list, cblk_count, st);
    return launches;
}
size_t tmk_scan_tile_desc_bytes(uint32_t) { return 16; }
#endif  // TMK_SCAN_ARM_A
