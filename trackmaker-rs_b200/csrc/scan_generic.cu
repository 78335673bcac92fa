// scan_generic.cu — preamble scan for ANY (line code, spl, preamble_bytes): sliding correlation
// with energy normalisation over every window of every stream, one pass over HBM.
//
// Replaces the per-sample loop of PhyDecoder::search_for_preamble (src/phy/decoder.rs:138-147)
// as a *filter*: it marks every window whose correlation may reach the threshold
// (corr_approx >= thr * TMK_CAND_RATIO) in a bitmap and summarises the window energies of each
// block of TMK_BLK windows as (min, max) plus the largest squared correlation of the block. The exact reference decision (AVX summation order,
// rolling energy) is re-established for the marked windows by the walk kernel (walk.cu).
//
// This translation unit is the general-shape kernel; scan_fast.cu holds the kernels specialised
// for the reference's shipped constants (spl = 3, 2 preamble bytes; src/utils/consts.rs:24-37).
#include "tmk_internal.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kSubBlks = 8;                      // blocks per CTA: one warp per block
constexpr int kSub = kSubBlks * TMK_BLK;         // windows per CTA (a scan tile is TMK_TILE_BLKS / kSubBlks of them)
constexpr int kSubPerTile = TMK_TILE_BLKS / kSubBlks;
static_assert(TMK_TILE_BLKS % kSubBlks == 0 && kSubBlks * 32 == kThreads, "sub-tiles of the generic scan");
constexpr int kSmemX = kSub + TMK_MAX_P + 8;

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


__global__ void __launch_bounds__(kThreads)
scan_generic_kernel(const float* __restrict__ samples, const TmkStreamDev* __restrict__ streams,
                    const uint32_t* __restrict__ tile_stream, uint32_t tile0, const TmkPattern* __restrict__ patterns,
                    float4* __restrict__ blk_info, uint32_t* __restrict__ cand_bits, uint32_t* __restrict__ cblk_list,
                    unsigned int* __restrict__ cblk_count) {
    __shared__ float xs[kSmemX];   // raw samples, then box sums of spl samples
    __shared__ float qs[kSmemX];   // box sums of spl squared samples
    __shared__ float lev[TMK_MAX_LEVELS];

    const uint32_t tile = tile0 + blockIdx.x / kSubPerTile;
    const TmkStreamDev sd = streams[tile_stream[tile]];
    const TmkPattern* pat = patterns + sd.pattern;
    if (pat->spl == 3 && pat->preamble_bytes == 2) return;  // scan_fast.cu owns the shipped shapes
    const uint32_t P = pat->P, L = pat->L, spl = pat->spl;
    const uint64_t n = sd.n_samples;
    const uint64_t nwin = n - P + 1;  // n >= P, otherwise the stream has no tiles
    const uint64_t w0 = uint64_t(tile - sd.tile_base) * TMK_TILE + uint64_t(blockIdx.x % kSubPerTile) * kSub;
    if (w0 >= nwin) return;  // a sub-tile past the end of the stream
    const float* __restrict__ x = samples + sd.offset;

    for (uint32_t l = threadIdx.x; l < L; l += kThreads) lev[l] = pat->levels[l];

    // stage the tile (+ P - 1 halo) in shared memory: coalesced float4 loads, zero fill past the end
    const uint32_t need = kSub + P - 1;
    for (uint32_t i = threadIdx.x * 4; i < need; i += kThreads * 4) {
        uint64_t g = w0 + i;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g + 3 < n) {
            v = __ldg(reinterpret_cast<const float4*>(x + g));
        } else {
            if (g < n) v.x = x[g];
            if (g + 1 < n) v.y = x[g + 1];
            if (g + 2 < n) v.z = x[g + 2];
        }
        *reinterpret_cast<float4*>(xs + i) = v;
    }
    for (uint32_t i = ((need + 3) & ~3u) + threadIdx.x; i < kSmemX; i += kThreads) xs[i] = 0.f;
    __syncthreads();

    // box sums over one level (spl samples) of x and x^2, in place via registers
    {
        float yb[(kSmemX + kThreads - 1) / kThreads], qb[(kSmemX + kThreads - 1) / kThreads];
        int c = 0;
        for (uint32_t i = threadIdx.x; i < need; i += kThreads, ++c) {
            float y = 0.f, q = 0.f;
            for (uint32_t k = 0; k < spl; ++k) {
                float v = (i + k < kSmemX) ? xs[i + k] : 0.f;
                y += v;
                q = fmaf(v, v, q);
            }
            yb[c] = y;
            qb[c] = q;
        }
        __syncthreads();
        c = 0;
        for (uint32_t i = threadIdx.x; i < need; i += kThreads, ++c) {
            xs[i] = yb[c];
            qs[i] = qb[c];
        }
    }
    __syncthreads();

    const float thr_c = sd.threshold * TMK_CAND_RATIO;
    const float kappa = thr_c * thr_c * float(P);
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bw0 = warp * TMK_BLK;  // this warp's block inside the tile
    float emin = __int_as_float(0x7f800000), emax = 0.f, dmax = 0.f;
    uint32_t myword = 0, ncand = 0;
#pragma unroll 1
    for (uint32_t i = 0; i < TMK_BLK_WORDS; ++i) {
        const uint32_t w = bw0 + i * 32 + lane;
        float D = 0.f, E = 0.f;
        for (uint32_t l = 0; l < L; ++l) {
            D = fmaf(lev[l], xs[w + l * spl], D);
            E += qs[w + l * spl];
        }
        const bool valid = (w0 + w) < nwin;
        const bool cand = valid && D > 0.f && D * D >= kappa * E;
        if (valid) {
            emin = fminf(emin, E);
            emax = max_nan(emax, E);
            if (D > 0.f) dmax = fmaxf(dmax, D * D / (float(P) * E));  // largest corr^2 of the block
        }
        const uint32_t word = __ballot_sync(0xffffffffu, cand);
        ncand += __popc(word);
        if (lane == i) myword = word;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        emin = fminf(emin, __shfl_xor_sync(0xffffffffu, emin, o));
        emax = max_nan(emax, __shfl_xor_sync(0xffffffffu, emax, o));
        dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    }
    const uint64_t blk_in_stream = w0 / TMK_BLK + warp;
    if (blk_in_stream * TMK_BLK < nwin) {
        const uint64_t gb = sd.blk_base + blk_in_stream;
        if (lane < TMK_BLK_WORDS) cand_bits[gb * TMK_BLK_WORDS + lane] = myword;
        if (lane == 0) {
            blk_info[gb] = make_float4(emin, emax, dmax, float(ncand));
            if (ncand && cblk_list) cblk_list[atomicAdd(cblk_count, 1u)] = uint32_t(gb);  // blocks with candidates
        }
    }
}

}  // namespace

int tmk_launch_tile_dir(const TmkStreamDev* streams, uint32_t n_streams, uint32_t n_tiles, uint32_t* tile_stream,
                        cudaStream_t st) {
    if (n_tiles == 0) return 0;
    tile_dir_kernel<<<(n_tiles + 255) / 256, 256, 0, st>>>(streams, n_streams, n_tiles, tile_stream);
    return 1;
}

int tmk_launch_scan_generic(const float* samples, const TmkStreamDev* streams, const uint32_t* tile_stream,
                            uint32_t tile0, uint32_t n_tiles, const TmkPattern* patterns, float4* blk_info, uint32_t* cand_bits,
                            uint32_t* cblk_list, unsigned int* cblk_count,
                            cudaStream_t st) {
    if (n_tiles == 0) return 0;
    scan_generic_kernel<<<n_tiles * kSubPerTile, kThreads, 0, st>>>(samples, streams, tile_stream, tile0, patterns, blk_info, cand_bits, cblk_list,
                                                            cblk_count);
    return 1;
}
