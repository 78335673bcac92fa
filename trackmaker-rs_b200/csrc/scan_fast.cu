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
constexpr int kWarps = 4;    // warps per CTA (each works on its own tiles)
constexpr unsigned kFull = 0xffffffffu;

template <int LC> struct Shape;
template <> struct Shape<TMK_LINE_4B5B> {
    static constexpr int L = 20, P = 60;
    static constexpr int kCopy = 180;  // floats copied per lane row (>= kS + P - 1, multiple of 4)
    static constexpr int kRow = 180;   // row pitch in floats: 720 B = 45 x 16 B
};
template <> struct Shape<TMK_LINE_MANCHESTER> {
    static constexpr int L = 32, P = 96;
    static constexpr int kCopy = 216;
    static constexpr int kRow = 220;   // 880 B = 55 x 16 B
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

// finish window n of the strip: candidate test and block statistics
template <bool FULL>
__device__ __forceinline__ void finish_window(int n, float D, float E, float kap, int nvalid, uint32_t* __restrict__ bits,
                                              uint32_t bit0, LaneOut& o) {
    const float dp = fmaxf(D, 0.0f);
    const float r = dp * dp * rcp_approx(E);  // P * corr^2 (NaN for silence: never a candidate)
    const bool ok = FULL || n < nvalid;
    if (FULL) {
        o.emin = fminf(o.emin, E);
        o.emax = max_nan(o.emax, E);
        o.rmax = fmaxf(o.rmax, r);
    } else if (ok) {
        o.emin = fminf(o.emin, E);
        o.emax = max_nan(o.emax, E);
        o.rmax = fmaxf(o.rmax, r);
    }
    if (r > kap && ok) {  // rare
        const uint32_t b = bit0 + uint32_t(n);
        atomicOr(bits + (b >> 5), 1u << (b & 31));
        o.ncand++;
    }
}

// ---------------------------------------------------------------- 4B5B: levels --++-++--++--+-++-++
// D(n) = -d[n] - y[n+12] + p2[n+15] - d[n+21] - p2[n+33] + y[n+39] - y[n+42] + p2[n+45] - y[n+51] + p2[n+54]
//   y[m] = x[m]+x[m+1]+x[m+2]   p2[m] = y[m]+y[m+3]   d[m] = p2[m]-p2[m+6]        (indices in samples)
template <bool FULL>
__device__ __forceinline__ void strip_4b5b(const float* __restrict__ row, float kap, int nvalid,
                                           uint32_t* __restrict__ bits, uint32_t bit0, LaneOut& o) {
    constexpr int P = 60;
    constexpr int M = kS + P - 3;  // level sums needed: y[0 .. M-1]
    float acc[64], yb[4], p2b[8], db[4], xr[8], xo[4];
    float e = 0.0f;
    {
        const float4 v = *reinterpret_cast<const float4*>(row);
        xr[0] = v.x; xr[1] = v.y; xr[2] = v.z; xr[3] = v.w;
        e = fmaf(v.x, v.x, e);
        e = fmaf(v.y, v.y, e);
    }
#pragma unroll
    for (int m = 0; m < M; ++m) {
        if (((m + 2) & 3) == 0) {  // x[m+2 .. m+5] become needed
            const float4 v = *reinterpret_cast<const float4*>(row + m + 2);
            xr[(m + 2) & 7] = v.x; xr[(m + 3) & 7] = v.y; xr[(m + 4) & 7] = v.z; xr[(m + 5) & 7] = v.w;
        }
        const float x2 = xr[(m + 2) & 7];
        const float y = (xr[m & 7] + xr[(m + 1) & 7]) + x2;
        e = fmaf(x2, x2, e);  // window energy: x[m+2] enters
        yb[m & 3] = y;
        float p2 = 0.f, d = 0.f;
        if (m >= 3) {
            p2 = yb[(m - 3) & 3] + y;  // p2[m-3]
            p2b[(m - 3) & 7] = p2;
        }
        if (m >= 9) {
            d = p2b[(m - 9) & 7] - p2;  // d[m-9]
            db[(m - 9) & 3] = d;
        }
        if (m - 12 >= 0 && m - 12 < kS) acc[(m - 12) & 63] = -db[(m - 12) & 3] - y;
        if (m - 18 >= 0 && m - 18 < kS) acc[(m - 18) & 63] += p2;
        if (m - 30 >= 0 && m - 30 < kS) acc[(m - 30) & 63] -= d;
        if (m - 36 >= 0 && m - 36 < kS) acc[(m - 36) & 63] -= p2;
        if (m - 39 >= 0 && m - 39 < kS) acc[(m - 39) & 63] += y;
        if (m - 42 >= 0 && m - 42 < kS) acc[(m - 42) & 63] -= y;
        if (m - 48 >= 0 && m - 48 < kS) acc[(m - 48) & 63] += p2;
        if (m - 51 >= 0 && m - 51 < kS) acc[(m - 51) & 63] -= y;
        if (m - 57 >= 0) {
            const int n = m - 57;  // window n is complete: x[n .. n+59] have entered e
            const float D = acc[n & 63] + p2;
            finish_window<FULL>(n, D, e, kap, nvalid, bits, bit0, o);
            if ((n & 3) == 0) {
                const float4 v = *reinterpret_cast<const float4*>(row + n);
                xo[0] = v.x; xo[1] = v.y; xo[2] = v.z; xo[3] = v.w;
            }
            const float xl = xo[n & 3];
            e = fmaf(-xl, xl, e);  // x[n] leaves
        }
    }
}

// ---------------------------------------------------------------- Manchester, bits 0011 0011 0101 1010
// D(n) = v[n] + v[n+24] + t[n+48] - t[n+72]
//   z[m] = y[m]-y[m+3]  w[m] = z[m]+z[m+6]  u[m] = z[m]-z[m+6]  v[m] = w[m]-w[m+12]  t[m] = u[m]+u[m+12]
template <bool FULL>
__device__ __forceinline__ void strip_manchester(const float* __restrict__ row, float kap, int nvalid,
                                                 uint32_t* __restrict__ bits, uint32_t bit0, LaneOut& o) {
    constexpr int P = 96;
    constexpr int M = kS + P - 3;
    float acc[128], yb[4], zb[8], wb[16], ub[16], xr[8], xo[4];
    float e = 0.0f;
    {
        const float4 v = *reinterpret_cast<const float4*>(row);
        xr[0] = v.x; xr[1] = v.y; xr[2] = v.z; xr[3] = v.w;
        e = fmaf(v.x, v.x, e);
        e = fmaf(v.y, v.y, e);
    }
#pragma unroll
    for (int m = 0; m < M; ++m) {
        if (((m + 2) & 3) == 0) {
            const float4 v = *reinterpret_cast<const float4*>(row + m + 2);
            xr[(m + 2) & 7] = v.x; xr[(m + 3) & 7] = v.y; xr[(m + 4) & 7] = v.z; xr[(m + 5) & 7] = v.w;
        }
        const float x2 = xr[(m + 2) & 7];
        const float y = (xr[m & 7] + xr[(m + 1) & 7]) + x2;
        e = fmaf(x2, x2, e);
        yb[m & 3] = y;
        float z = 0.f, w = 0.f, u = 0.f, vv = 0.f, tt = 0.f;
        if (m >= 3) {
            z = yb[(m - 3) & 3] - y;  // z[m-3]
            zb[(m - 3) & 7] = z;
        }
        if (m >= 9) {
            w = zb[(m - 9) & 7] + z;  // w[m-9], u[m-9]
            u = zb[(m - 9) & 7] - z;
            wb[(m - 9) & 15] = w;
            ub[(m - 9) & 15] = u;
        }
        if (m >= 21) {
            vv = wb[(m - 21) & 15] - w;  // v[m-21], t[m-21]
            tt = ub[(m - 21) & 15] + u;
        }
        if (m - 21 >= 0 && m - 21 < kS) acc[(m - 21) & 127] = vv;
        if (m - 45 >= 0 && m - 45 < kS) acc[(m - 45) & 127] += vv;
        if (m - 69 >= 0 && m - 69 < kS) acc[(m - 69) & 127] += tt;
        if (m - 93 >= 0) {
            const int n = m - 93;
            const float D = acc[n & 127] - tt;
            finish_window<FULL>(n, D, e, kap, nvalid, bits, bit0, o);
            if ((n & 3) == 0) {
                const float4 v = *reinterpret_cast<const float4*>(row + n);
                xo[0] = v.x; xo[1] = v.y; xo[2] = v.z; xo[3] = v.w;
            }
            const float xl = xo[n & 3];
            e = fmaf(-xl, xl, e);
        }
    }
}

template <int LC>
__global__ void __launch_bounds__(kWarps * 32, 2)
scan_fast_kernel(const float* __restrict__ samples, uint64_t total_samples, const TmkStreamDev* __restrict__ streams,
                 const uint32_t* __restrict__ tile_stream, uint32_t n_tiles, const TmkPattern* __restrict__ patterns,
                 float4* __restrict__ blk_info, uint32_t* __restrict__ cand_bits) {
    using Sh = Shape<LC>;
    extern __shared__ __align__(128) float smem[];
    __shared__ __align__(8) uint64_t mbar[kWarps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* row = smem + size_t(warp * 32 + lane) * Sh::kRow;
    uint64_t* bar = &mbar[warp];
    if (lane == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    uint32_t parity = 0;
    for (uint32_t tile = blockIdx.x * kWarps + warp; tile < n_tiles; tile += gridDim.x * kWarps) {
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
        if ((lane & 3) == 0 && w0 + uint64_t(lane >> 2) * TMK_BLK < nwin)
            blk_info[blk0 + (lane >> 2)] = make_float4(o.emin, o.emax, o.rmax * (1.0f / float(Sh::P)), ncf);
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
                 uint32_t n_tiles, const TmkPattern* patterns, float4* blk_info, uint32_t* cand_bits, cudaStream_t st) {
    static int grid_cap = 0;
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
    scan_fast_kernel<LC><<<grid, kWarps * 32, smem, st>>>(samples, total_samples, streams, tile_stream, n_tiles,
                                                        patterns, blk_info, cand_bits);
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
                    const uint32_t* tile_stream, uint32_t n_tiles, const TmkPattern* patterns, uint32_t kinds,
                    float4* blk_info, uint32_t* cand_bits, cudaStream_t st) {
    if (n_tiles == 0) return 0;
    int launches = 0;
    if (kinds & 1u) {
        launch_fast<TMK_LINE_4B5B>(samples, total_samples, streams, tile_stream, n_tiles, patterns, blk_info, cand_bits, st);
        ++launches;
    }
    if (kinds & 2u) {
        launch_fast<TMK_LINE_MANCHESTER>(samples, total_samples, streams, tile_stream, n_tiles, patterns, blk_info,
                                         cand_bits, st);
        ++launches;
    }
    if (kinds & 4u) launches += tmk_launch_scan_generic(samples, streams, tile_stream, n_tiles, patterns, blk_info, cand_bits, st);
    return launches;
}
