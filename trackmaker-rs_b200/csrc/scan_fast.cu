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
#include "tmk_internal.cuh"

namespace {

constexpr int kS = 120;      // windows per lane strip
constexpr int kMaxWarps = 4; // warps per CTA (each works on its own tiles): Shape<LC>::kWarps <= kMaxWarps
constexpr unsigned kFull = 0xffffffffu;

template <int LC> struct Shape;
template <> struct Shape<TMK_LINE_4B5B> {
    static constexpr int L = 20, P = 60;
    static constexpr int kCopy = 184;  // floats copied per lane row (3 x 60 steps read x[0 .. 183]), multiple of 4
    static constexpr int kRow = 188;   // row pitch in floats: 752 B = 47 x 16 B (odd: conflict-free LDS.128)
    static constexpr int kWarps = 4;   // 96 KB of rows per CTA, 2 CTAs per SM
};
template <> struct Shape<TMK_LINE_MANCHESTER> {
    static constexpr int L = 32, P = 96;
    static constexpr int kCopy = 232;  // 3 x 76 steps read x[0 .. 231]
    static constexpr int kRow = 236;   // 944 B = 59 x 16 B
    static constexpr int kWarps = 3;   // 91 KB of rows per CTA, 2 CTAs per SM
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

// finish window n of the strip (branch free): block statistics and the candidate predicate.
// `valid` masks the windows that do not belong to the strip (warm-up, past the strip / the stream).
__device__ __forceinline__ bool finish_window(float D, float E, float kap, bool valid, LaneOut& o) {
    const float dp = fmaxf(D, 0.0f);
    const float Ev = valid ? E : __int_as_float(0x7fc00000);  // NaN: ignored by fminf / fmaxf, never a candidate
    const float r = dp * dp * rcp_approx(Ev);                  // P * corr^2 (NaN for silence: never a candidate)
    o.emin = fminf(o.emin, Ev);
    o.emax = max_nan(o.emax, valid ? E : 0.0f);                // a NaN SAMPLE must stay visible in max E
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

// Which loop iterations k (of NIT) make step j finish a window of the strip: n = U * k + j - LAT in [0, hi).
// j is a compile-time constant after unrolling, so this folds to an OR of at most NIT iteration flags.
template <int U, int LAT, int NIT>
__device__ __forceinline__ bool step_valid(int j, int hi, const bool (&is)[NIT]) {
    bool v = false;
#pragma unroll
    for (int kk = 0; kk < NIT; ++kk) {
        const int n = U * kk + j - LAT;
        if (n >= 0 && n < hi) v = v || is[kk];
    }
    return v;
}

// The strips are ROLLED loops of NIT iterations x U steps: the unrolled body is the same ~3 K instructions
// for every warp and stays resident in the instruction cache (the fully unrolled 177 / 213-step form was
// 148 KB of straight-line code streamed by every warp). All ring indices are static because the ring sizes
// divide U; warm-up steps run the same code on zeroed rings and are masked when a window is finished.

// ---------------------------------------------------------------- 4B5B: levels --++-++--++--+-++-++
// D(n) = -d[n] - y[n+12] + p2[n+15] - d[n+21] - p2[n+33] + y[n+39] - y[n+42] + p2[n+45] - y[n+51] + p2[n+54]
//   y[m] = x[m]+x[m+1]+x[m+2]   p2[m] = y[m]+y[m+3]   d[m] = p2[m]-p2[m+6]        (indices in samples)
template <bool FULL>
__device__ __forceinline__ void strip_4b5b(const float* __restrict__ row, float kap, int nvalid,
                                           uint32_t* __restrict__ bits, uint32_t bit0, LaneOut& o) {
    constexpr int U = 60, LAT = 57, NIT = 3;  // steps per iteration, steps from a window's start to its finish
    static_assert(U * NIT >= kS + LAT && U % 12 == 0 && U <= 64, "strip steps");
    float acc[60], yb[4], p2b[12], db[4], xr[12], xo[4];
#pragma unroll
    for (int i = 0; i < 60; ++i) acc[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 12; ++i) p2b[i] = xr[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) yb[i] = db[i] = xo[i] = 0.f;
    float e = 0.0f;
    {
        const float4 v = *reinterpret_cast<const float4*>(row);
        xr[0] = v.x; xr[1] = v.y; xr[2] = v.z; xr[3] = v.w;
        e = fmaf(v.x, v.x, e);
        e = fmaf(v.y, v.y, e);
    }
#pragma unroll 1
    for (int k = 0; k < NIT; ++k) {
        const float* __restrict__ rk = row + U * k;
        const bool is[NIT] = {k == 0, k == 1, k == 2};
        uint64_t cm = 0;
#pragma unroll
        for (int j = 0; j < U; ++j) {  // m = U * k + j
            if (((j + 2) & 3) == 0) {  // x[m+2 .. m+5] become needed
                const float4 v = *reinterpret_cast<const float4*>(rk + j + 2);
                xr[rix<12>(j + 2)] = v.x; xr[rix<12>(j + 3)] = v.y; xr[rix<12>(j + 4)] = v.z; xr[rix<12>(j + 5)] = v.w;
            }
            const float x2 = xr[rix<12>(j + 2)];
            const float y = (xr[rix<12>(j)] + xr[rix<12>(j + 1)]) + x2;
            e = fmaf(x2, x2, e);  // window energy: x[m+2] enters
            const float p2 = yb[rix<4>(j - 3)] + y;    // p2[m-3]
            yb[rix<4>(j)] = y;
            const float d = p2b[rix<12>(j - 9)] - p2;  // d[m-9]
            p2b[rix<12>(j - 3)] = p2;
            acc[rix<60>(j - 12)] = -db[rix<4>(j - 12)] - y;
            db[rix<4>(j - 9)] = d;
            acc[rix<60>(j - 18)] += p2;
            acc[rix<60>(j - 30)] -= d;
            acc[rix<60>(j - 36)] -= p2;
            acc[rix<60>(j - 39)] += y;
            acc[rix<60>(j - 42)] -= y;
            acc[rix<60>(j - 48)] += p2;
            acc[rix<60>(j - 51)] -= y;
            // window n = m - 57 is complete: x[n .. n+59] have entered e
            const float D = acc[rix<60>(j - LAT)] + p2;
            bool valid = step_valid<U, LAT, NIT>(j, kS, is);
            if (!FULL) valid = valid && (U * k + j - LAT) < nvalid;
            if (finish_window(D, e, kap, valid, o)) cm |= 1ull << j;
            if (((j - LAT) & 3) == 0) {  // x[n .. n+3] (row start for the warm-up steps: unused)
                const int nn = U * k + j - LAT;
                const float4 v = *reinterpret_cast<const float4*>(row + (nn >= 0 ? nn : 0));
                xo[0] = v.x; xo[1] = v.y; xo[2] = v.z; xo[3] = v.w;
            }
            const float xl = step_valid<U, LAT, NIT>(j, 1 << 20, is) ? xo[(j - LAT) & 3] : 0.0f;
            e = fmaf(-xl, xl, e);  // x[n] leaves
        }
        if (cm) flush_cands<U, LAT>(cm, 0, k, FULL ? kS : nvalid, bits, bit0, o);
    }
}

// ---------------------------------------------------------------- Manchester, bits 0011 0011 0101 1010
// D(n) = v[n] + v[n+24] + t[n+48] - t[n+72]
//   z[m] = y[m]-y[m+3]  w[m] = z[m]+z[m+6]  u[m] = z[m]-z[m+6]  v[m] = w[m]-w[m+12]  t[m] = u[m]+u[m+12]
template <bool FULL>
__device__ __forceinline__ void strip_manchester(const float* __restrict__ row, float kap, int nvalid,
                                                 uint32_t* __restrict__ bits, uint32_t bit0, LaneOut& o) {
    constexpr int U = 76, LAT = 93, NIT = 3;
    static_assert(U * NIT >= kS + LAT && U % 19 == 0 && U % 4 == 0 && U <= 128, "strip steps");
    float acc[76], yb[4], zb[19], wb[19], ub[19], xr[19], xo[4];
#pragma unroll
    for (int i = 0; i < 76; ++i) acc[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 19; ++i) zb[i] = wb[i] = ub[i] = xr[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) yb[i] = xo[i] = 0.f;
    float e = 0.0f;
    {
        const float4 v = *reinterpret_cast<const float4*>(row);
        xr[0] = v.x; xr[1] = v.y; xr[2] = v.z; xr[3] = v.w;
        e = fmaf(v.x, v.x, e);
        e = fmaf(v.y, v.y, e);
    }
#pragma unroll 1
    for (int k = 0; k < NIT; ++k) {
        const float* __restrict__ rk = row + U * k;
        const bool is[NIT] = {k == 0, k == 1, k == 2};
        uint64_t cm0 = 0, cm1 = 0;
#pragma unroll
        for (int j = 0; j < U; ++j) {  // m = U * k + j
            if (((j + 2) & 3) == 0) {
                const float4 v = *reinterpret_cast<const float4*>(rk + j + 2);
                xr[rix<19>(j + 2)] = v.x; xr[rix<19>(j + 3)] = v.y; xr[rix<19>(j + 4)] = v.z; xr[rix<19>(j + 5)] = v.w;
            }
            const float x2 = xr[rix<19>(j + 2)];
            const float y = (xr[rix<19>(j)] + xr[rix<19>(j + 1)]) + x2;
            e = fmaf(x2, x2, e);
            const float z = yb[rix<4>(j - 3)] - y;  // z[m-3]
            yb[rix<4>(j)] = y;
            const float zo = zb[rix<19>(j - 9)];
            const float w = zo + z, u = zo - z;      // w[m-9], u[m-9]
            zb[rix<19>(j - 3)] = z;
            const float vv = wb[rix<19>(j - 21)] - w;  // v[m-21], t[m-21]
            const float tt = ub[rix<19>(j - 21)] + u;
            wb[rix<19>(j - 9)] = w;
            ub[rix<19>(j - 9)] = u;
            acc[rix<76>(j - 21)] = vv;
            acc[rix<76>(j - 45)] += vv;
            acc[rix<76>(j - 69)] += tt;
            const float D = acc[rix<76>(j - LAT)] - tt;  // window n = m - 93
            bool valid = step_valid<U, LAT, NIT>(j, kS, is);
            if (!FULL) valid = valid && (U * k + j - LAT) < nvalid;
            if (finish_window(D, e, kap, valid, o)) {
                if (j < 64) cm0 |= 1ull << (j & 63);
                else cm1 |= 1ull << (j & 63);
            }
            if (((j - LAT) & 3) == 0) {
                const int nn = U * k + j - LAT;
                const float4 v = *reinterpret_cast<const float4*>(row + (nn >= 0 ? nn : 0));
                xo[0] = v.x; xo[1] = v.y; xo[2] = v.z; xo[3] = v.w;
            }
            const float xl = step_valid<U, LAT, NIT>(j, 1 << 20, is) ? xo[(j - LAT) & 3] : 0.0f;
            e = fmaf(-xl, xl, e);
        }
        if (cm0 | cm1) flush_cands<U, LAT>(cm0, cm1, k, FULL ? kS : nvalid, bits, bit0, o);
    }
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

// stream that owns each scan tile (streams are sorted by tile_base)
__global__ void tile_dir_kernel(const TmkStreamDev* __restrict__ streams, uint32_t n_streams, uint32_t n_tiles,
                                uint32_t* __restrict__ tile_stream) {
    const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= n_tiles) return;
    uint32_t lo = 0, hi = n_streams - 1;
    while (lo < hi) {
        const uint32_t mid = (lo + hi + 1) >> 1;
        if (streams[mid].tile_base <= tile) lo = mid; else hi = mid - 1;
    }
    // streams without tiles share their tile_base with the next one: lo is the last of them, and the
    // only one that has tiles, because the search keeps the largest index with tile_base <= tile
    tile_stream[tile] = lo;
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

int tmk_launch_tile_dir(const TmkStreamDev* streams, uint32_t n_streams, uint32_t n_tiles, uint32_t* tile_stream,
                        cudaStream_t st) {
    if (n_tiles == 0) return 0;
    tile_dir_kernel<<<(n_tiles + 255) / 256, 256, 0, st>>>(streams, n_streams, n_tiles, tile_stream);
    return 1;
}

// kinds: bit 0 = some stream is 4B5B/3/2, bit 1 = Manchester/3/2, bit 2 = any other shape (generic kernel)
int tmk_launch_scan(const float* samples, uint64_t total_samples, const TmkStreamDev* streams, uint32_t n_streams,
                    const uint32_t* tile_stream, uint32_t tile0, uint32_t n_tiles, const TmkPattern* patterns,
                    uint32_t kinds, float4* blk_info, uint32_t* cand_bits, uint32_t* cblk_list, unsigned int* cblk_count,
                    cudaStream_t st) {
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
