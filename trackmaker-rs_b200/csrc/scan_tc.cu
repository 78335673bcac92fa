// scan_tc.cu — preamble scan for the reference's shipped constants (SAMPLES_PER_LEVEL = 3,
// PREAMBLE_PATTERN_BYTES = 2; src/utils/consts.rs:30-34) on the 5th-generation tensor cores: the hot kernel of
// the engine, one pass over HBM, 4 B per sample.
//
// It replaces the per-sample loop of PhyDecoder::search_for_preamble (src/phy/decoder.rs:138-147: P multiply-adds
// + sqrt + divide per sample) as a FILTER (contract in scan_generic.cu): a candidate bit per window and
// (min E, max E, max corr^2, #candidates) per block of 256 windows. The exact decision is taken later, only for
// marked windows (walk.cu), so reduced precision here cannot change a result — which is what lets the sliding
// +-1 correlation run as a banded-Toeplitz GEMM in TF32:
//
//   the capture is read as rows of 64 samples; a tile is 128 rows (8192 windows, 32 blocks).
//   D[row r][offset n] = sum_t x[64 r + n + t] p[t]  =  sum_h  X_h (128 x 64)  .  T_h^T (64 x 64)
//   X_h[r][k] = x[64 (r + h) + k]  (the tile's rows shifted by h rows),  T_h[n][k] = p[64 h + k - n]  (constant)
//
//   TMA      one 4-D tensor-map box per tile (128-byte rows, 128-byte swizzle) lands the rows in shared memory
//            ALREADY in the canonical K-major SWIZZLE_128B layout of tcgen05, with the MMA row i = 8 a + b holding
//            capture row 16 b + a: "the next capture row" is then "+1 in a", a constant +SBO on the operand
//            address, so X_1 (X_2) is the same buffer with a shifted descriptor and every sample is staged once
//            (the one or two row groups of the next rows are the only duplicates, 6 %). (A first version used the
//            unswizzled core-matrix layout: its 16-byte box rows made 2176 TMA requests per tile and the kernel
//            request-bound at 2.8 ms; 128-byte rows are 272.)
//   UMMA     tcgen05.mma kind::tf32, M = 128, N = 64, K = 8 per instruction: 16 (4B5B, K = 128) or 20
//            (Manchester, K = 160) instructions per tile, issued by one thread, accumulating in TMEM
//            (2 accumulator stages of 64 columns).
//   epilogue four warps read their TMEM lane quarter (tcgen05.ld 32x32b: one row = 64 consecutive windows per
//            thread), roll the window energy out of the staged samples in fp32 (one FFMA in, one out), test
//            D+^2 / E > P (0.85 thr / 0.9)^2 with rcp.approx, and reduce the block statistics. Candidate words are
//            plain stores (a row owns two whole bitmap words): no atomics, no memset of the bitmap.
//   The producer / MMA / epilogue roles are decoupled by mbarriers (4 shared-memory stages, 2 TMEM stages);
//   CTAs are persistent, one per SM.
//
// TF32 keeps 10 mantissa bits of every sample (the +-1 pattern is exact): |D_tf32 - D| <= 2^-10 sum |x| <=
// 2^-10 sqrt(P E), i.e. at most 0.1 % of the normalised correlation, inside the 0.9 % the walk leaves for the
// filter's own error (kTight, walk.cu); max corr^2 is reported with that margin added.
#ifndef TMK_SCAN_ARM_A
#include <cuda.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "tmk_internal.cuh"

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kRowLen = 64;                    // samples per row = window offsets per row (N)
constexpr int kRows = 128;                     // rows per tile (M)
static_assert(kRows * kRowLen == TMK_TILE && TMK_BLK == 4 * kRowLen && TMK_BLK_WORDS == 8, "tile geometry");
constexpr int kTeams = 4;                      // epilogue teams of four warps (one per TMEM lane quarter); team = iteration % kTeams
constexpr int kStages = kTeams;                // shared-memory stages (sample tiles in flight): one per team
constexpr int kAccStages = kTeams;             // TMEM accumulator stages: one per team
constexpr int kEpiWarps = 4 * kTeams;
constexpr int kThreads = 32 * (2 + kEpiWarps);
constexpr uint32_t kGroupBytes = 2048;         // one row group `a`: two 128B-swizzle atoms (sample columns 0..31 | 32..63) of 8 rows x 128 B
constexpr uint32_t kAtomBytes = 1024;
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(kRowLen >> 3) << 17) | (uint32_t(kRows >> 4) << 24);

template <int LC> struct Tc;
// preamble levels of LineCode::generate_preamble(2) (line_coding.rs:13-21: 0x33 0x5A line coded), one char per level
template <> struct Tc<TMK_LINE_4B5B> {
    static constexpr int P = 60, kHalves = 2, kGroups = 17;  // K = 128: rows r, r + 1
    static constexpr const char* kLevels = "--++-++--++--+-++-++";
};
template <> struct Tc<TMK_LINE_MANCHESTER> {
    static constexpr int P = 96, kHalves = 3, kGroups = 18;  // K = 160: rows r, r + 1, first half of r + 2
    static constexpr const char* kLevels = "+-+--+-++-+--+-++--++--+-++--++-";
};
constexpr int kMaxMaps = 20;          // one tensor map per window of kMapWin samples of the batch buffer (int32 box coordinate)
constexpr uint64_t kMapWin = (1ull << 31) - (1ull << 16);
struct TcMaps {
    CUtensorMap m[kMaxMaps];
};

// what a role needs to know about a scan tile: one 32-byte load per tile instead of the chain
// tile_stream -> stream -> pattern
struct alignas(16) TileDesc {
    uint64_t start;   // first sample of the tile in the batch buffer
    uint64_t nwin;    // windows of the stream from the tile's first one on (0: a tile of another kernel)
    uint64_t blk0;    // global index of the tile's first block
    float thr;        // threshold of the stream
    uint32_t kind;    // 1 = 4B5B / 3 / 2, 2 = Manchester / 3 / 2, 0 = any other shape
};
static_assert(sizeof(TileDesc) == 32, "tile descriptor");

// dir_out (nullable): the kernel also finds the stream that owns each tile (what tile_dir_kernel does: streams are sorted by
// tile_base) and writes the directory — one launch less in front of the scan of a one-chunk batch
__global__ void tile_desc_kernel(const TmkStreamDev* __restrict__ streams, uint32_t n_streams, const uint32_t* __restrict__ tile_stream,
                                 uint32_t* __restrict__ dir_out, const TmkPattern* __restrict__ patterns, uint32_t tile0,
                                 uint32_t n_tiles, TileDesc* __restrict__ out) {
    // (small batches: the tile bases of all streams are fetched once, in parallel, and searched in shared memory — eight
    //  dependent trips to L2 per tile were most of this kernel's 7 us in front of a links poll)
    constexpr uint32_t kDirSmem = 2048;
    __shared__ uint32_t tb[kDirSmem];
    const bool in_smem = dir_out != nullptr && n_streams <= kDirSmem;
    if (in_smem) {
        for (uint32_t i = threadIdx.x; i < n_streams; i += blockDim.x) tb[i] = streams[i].tile_base;
        __syncthreads();
    }
    const uint32_t tl = blockIdx.x * blockDim.x + threadIdx.x;
    if (tl >= n_tiles) return;
    const uint32_t tile = tile0 + tl;
    uint32_t owner;
    if (dir_out != nullptr) {
        uint32_t lo = 0, hi = n_streams - 1;
        while (lo < hi) {
            const uint32_t mid = (lo + hi + 1) >> 1;
            if ((in_smem ? tb[mid] : streams[mid].tile_base) <= tile) lo = mid; else hi = mid - 1;
        }
        owner = lo;
        dir_out[tile] = lo;
    } else {
        owner = tile_stream[tile];
    }
    const TmkStreamDev sd = streams[owner];
    const TmkPattern* pat = patterns + sd.pattern;
    const uint64_t w0 = uint64_t(tile - sd.tile_base) * TMK_TILE;
    TileDesc d;
    d.start = sd.offset + w0;
    d.nwin = sd.n_samples - pat->P + 1 - w0;
    d.blk0 = sd.blk_base + w0 / TMK_BLK;
    d.thr = sd.threshold;
    d.kind = (pat->spl == 3 && pat->preamble_bytes == 2) ? (pat->line_code == TMK_LINE_4B5B ? 1u : 2u) : 0u;
    out[tl] = d;
}

struct Smem {
    // (dynamic shared memory, 1024-byte aligned) [stages x kGroups x 2048 | kHalves x 16 KB of T_h | ...]
    uint64_t full[kStages], empty[kStages], tfull[kAccStages], tempty[kAccStages];
    uint32_t tmem_base;
    uint32_t pad_;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
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
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// K-major, 128-byte swizzle: rows of 128 B (32 TF32 values of K), atoms of 8 rows (1024 B, 1024-byte aligned), the
// 16-byte chunk j of row r stored at chunk j ^ (r % 8); SBO = bytes between 8-row groups; an instruction (K = 8)
// starts 32 B further along the row (cute::UMMA::SmemDescriptor: start >> 4 at [0,14), LBO >> 4 at [16,30) — unused
// for swizzled K-major, 1 —, SBO >> 4 at [32,46), version 1 at [46,48), layout type 2 = SWIZZLE_128B at [61,64))
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t sbo) {
    return uint64_t((saddr >> 4) & 0x3fffu) | (1ull << 16) | (uint64_t((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// the same with the descriptors as (low word, high word): per instruction only the low words change, by a constant
// (see ask_tc.cu: 6 SASS instructions per MMA in the one issuing thread instead of 13 and a branch)
__device__ __forceinline__ void umma_tf32_lohi(uint32_t tmem_d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "mov.b64 da, {%1, %2};\n"
        "mov.b64 db, {%3, %4};\n"
        "setp.ne.b32 p, %6, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %5, p;\n"
        "}\n" ::"r"(tmem_d), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ uint32_t elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred P;\n"
        "elect.sync _|P, 0xffffffff;\n"
        "selp.b32 %0, 1, 0, P;\n"
        "}\n" : "=r"(pred));
    return pred;
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&r)[32]) {
    uint32_t u[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]),
          "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15]), "=r"(u[16]),
          "=r"(u[17]), "=r"(u[18]), "=r"(u[19]), "=r"(u[20]), "=r"(u[21]), "=r"(u[22]), "=r"(u[23]), "=r"(u[24]),
          "=r"(u[25]), "=r"(u[26]), "=r"(u[27]), "=r"(u[28]), "=r"(u[29]), "=r"(u[30]), "=r"(u[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = __uint_as_float(u[i]);
}

struct RowStat {
    float emin, emax, rmax;
    uint32_t ncand;
};

// one window of the epilogue. FULL: every window of the tile exists (no masking).
template <bool FULL>
__device__ __forceinline__ bool tc_window(float Dv, float e, float kap, bool valid, RowStat& o) {
    const float dp = fmaxf(Dv, 0.0f);
    if (FULL) {
        const float r = dp * dp * rcp_approx(e);     // P * corr^2 (NaN for silence: never a candidate)
        o.emin = fminf(o.emin, e);
        o.emax = max_nan(o.emax, e);                 // a NaN energy stays visible in max E
        o.rmax = fmaxf(o.rmax, r);
        return r > kap;
    }
    const float Ev = valid ? e : __int_as_float(0x7fc00000);  // NaN: ignored by fminf / fmaxf, never a candidate
    const float r = dp * dp * rcp_approx(Ev);
    o.emin = fminf(o.emin, Ev);
    o.emax = max_nan(o.emax, valid ? e : 0.0f);
    o.rmax = fmaxf(o.rmax, r);
    return r > kap;
}

// The 64 windows of one row: dot products from TMEM (two loads of 32 columns), the window energy rolled out of the
// staged samples, candidate mask and row statistics. `row`: the thread's row in the stage (un-swizzle with b).
template <int P, bool FULL>
__device__ __forceinline__ void epi_row(const uint8_t* __restrict__ row, int b, uint32_t taddr, float kap, int nvalid,
                                        uint64_t* tfull, uint32_t tphase, uint64_t* tempty, int lane, float* __restrict__ dbg,
                                        RowStat& o, uint32_t (&mask)[2]) {
    auto chunk = [&](int c) -> float4 {  // samples 4 c .. 4 c + 3 counted from the row start
        return *reinterpret_cast<const float4*>(row + (c >> 4) * kGroupBytes + ((c >> 3) & 1) * kAtomBytes + (((c & 7) ^ b) << 4));
    };
    float e = 0.0f;  // energy of the row's first window
#pragma unroll
    for (int c = 0; c < P / 4; ++c) {
        const float4 v = chunk(c);
        e = fmaf(v.x, v.x, e);
        e = fmaf(v.y, v.y, e);
        e = fmaf(v.z, v.z, e);
        e = fmaf(v.w, v.w, e);
    }
    // (the seed energy needs the samples only: it runs while the tile's MMAs are still in flight — a fresh profile had 38 % of
    //  the kernel's samples on this wait when it stood in front of the seed sum)
    mbar_wait(tfull, tphase);  // the dot products are in TMEM
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        float D[32];
        tmem_ld32(taddr + 32 * half, D);
        if (half == 1) {  // both halves are in registers: the accumulator stage is free again
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty);
        }
        // a NaN / inf sample anywhere in the rows of this MMA row turns ALL its dot products into NaN (0 x NaN in
        // the banded operand): one test per row keeps it visible in max E, like a NaN energy
        if (half == 0 && nvalid > 0 && !(fabsf(D[0]) <= 3.0e38f)) o.emax = __int_as_float(0x7fc00000);
        if (dbg != nullptr)
            for (int j = 0; j < 32; ++j)
                if (32 * half + j < nvalid) dbg[32 * half + j] = D[j];
        uint32_t mk = 0;
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
            const float4 xo = chunk(8 * half + c4), xi = chunk(8 * half + c4 + P / 4);
            const float xov[4] = {xo.x, xo.y, xo.z, xo.w}, xiv[4] = {xi.x, xi.y, xi.z, xi.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = 4 * c4 + u;
                if (tc_window<FULL>(D[j], e, kap, 32 * half + j < nvalid, o)) mk |= 1u << j;
                e = fmaf(xiv[u], xiv[u], e);    // x[n + P] enters
                e = fmaf(-xov[u], xov[u], e);   // x[n] leaves
            }
        }
        mask[half] = mk;
        o.ncand += __popc(mk);
    }
}

template <int LC>
__global__ void __launch_bounds__(kThreads, 1)
scan_tc_kernel(const __grid_constant__ TcMaps maps, int n_maps, const float* __restrict__ samples, uint64_t total_samples,
               const TileDesc* __restrict__ tdesc, uint32_t n_tiles, float4* __restrict__ blk_info,
               uint32_t* __restrict__ cand_bits, uint32_t* __restrict__ cblk_list, unsigned int* __restrict__ cblk_count,
               float* __restrict__ dbg_dot) {
    using S = Tc<LC>;
    constexpr int P = S::P;
    constexpr uint32_t kKind = LC == TMK_LINE_4B5B ? 1u : 2u;
    constexpr uint32_t kStageBytes = S::kGroups * kGroupBytes;
    constexpr uint32_t kTBytes = 16384;  // one T_h: column blocks c = 0, 1, each 8 atoms (64 rows n) of 1024 B
    extern __shared__ uint8_t smem_dyn[];
    uint8_t* smem_raw = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    uint8_t* stage0 = smem_raw;
    uint8_t* tmat = smem_raw + kStages * kStageBytes;
    Smem& sm = *reinterpret_cast<Smem*>(tmat + S::kHalves * kTBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // ---- one-time setup: barriers, TMEM, the Toeplitz operands T_h[n][k] = p[64 h + k - n]
    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&sm.full[i], 1);
            mbar_init(&sm.empty[i], 4);   // the four warps of the team that consumed the stage
        }
        for (int i = 0; i < kAccStages; ++i) {
            mbar_init(&sm.tfull[i], 1);
            mbar_init(&sm.tempty[i], 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                     "n"(kAccStages * kRowLen) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (uint32_t i = threadIdx.x; i < uint32_t(S::kHalves) * 64 * 64; i += kThreads) {
        const uint32_t h = i >> 12, n = (i >> 6) & 63, k = i & 63;
        const int t = int(64 * h + k) - int(n);
        const float v = (t >= 0 && t < P) ? (S::kLevels[t / 3] == '+' ? 1.0f : -1.0f) : 0.0f;
        const uint32_t c = k >> 5, kc = (k >> 2) & 7;
        *reinterpret_cast<float*>(tmat + h * kTBytes + c * 8192 + (n >> 3) * kAtomBytes + (n & 7) * 128 + ((kc ^ (n & 7)) << 4) + (k & 3) * 4) = v;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // T_h (generic stores) -> tensor-core reads (async proxy)
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = sm.tmem_base;

    // Every role walks the same sequence of tiles (tl = blockIdx.x, + gridDim.x, ...); `it` counts the tiles of THIS
    // kernel's line code among them. The descriptor of the next tile is loaded while the current one is worked on.
    auto load_desc = [&](uint32_t tl) -> TileDesc {
        TileDesc d;
        if (tl < n_tiles) {
            const uint4* p = reinterpret_cast<const uint4*>(tdesc + tl);
            const uint4 lo = __ldg(p), hi = __ldg(p + 1);
            d.start = uint64_t(lo.x) | (uint64_t(lo.y) << 32);
            d.nwin = uint64_t(lo.z) | (uint64_t(lo.w) << 32);
            d.blk0 = uint64_t(hi.x) | (uint64_t(hi.y) << 32);
            d.thr = __uint_as_float(hi.z);
            d.kind = hi.w;
        } else {
            d.start = d.nwin = d.blk0 = 0;
            d.thr = 0.f;
            d.kind = 0;
        }
        return d;
    };

    // Iteration j of a CTA is tile tl = blockIdx.x + j gridDim.x; team j % kTeams owns it, together with shared-memory
    // stage and TMEM stage j % kTeams. Tiles of another kernel's line code are skipped by everybody; cnt[team] counts
    // the tiles a team has processed (the phase of its barriers).
    if (warp == 0) {
        // ================================================================ producer: sample tiles -> shared memory
        uint32_t cnt[kTeams] = {};
        uint32_t j = 0;
        TileDesc nxt = load_desc(blockIdx.x);
        for (uint32_t tl = blockIdx.x; tl < n_tiles; tl += gridDim.x, ++j) {
            const TileDesc td = nxt;
            nxt = load_desc(tl + gridDim.x);
            if (td.kind != kKind) continue;
            const uint32_t s = j % kTeams;
            uint32_t ph = 0;
#pragma unroll
            for (int k = 0; k < kTeams; ++k)
                if (uint32_t(k) == s) ph = cnt[k]++ & 1u;
            mbar_wait(&sm.empty[s], ph ^ 1u);
            uint8_t* dst = stage0 + s * kStageBytes;
            const uint64_t start = td.start;
            const uint64_t span = uint64_t(16 * 7 + S::kGroups) * kRowLen;  // samples the box reaches
            const uint32_t mi = uint32_t(start / kMapWin);
            const bool tma = mi < uint32_t(n_maps) && start + span <= total_samples;
            if (tma) {
                if (lane == 0) {
                    mbar_expect_tx(&sm.full[s], kStageBytes);
                    const int32_t c0 = int32_t(start - uint64_t(mi) * kMapWin);  // in samples
                    asm volatile(
                        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(&maps.m[mi])), "r"(smem_u32(&sm.full[s])), "r"(c0), "r"(0),
                        "r"(0), "r"(0) : "memory");
                }
            } else {
                // the last tiles of the buffer (the box would run past its end) and unsupported maps: the warp fills
                // the same layout itself, zeros past the end
                for (uint32_t g = lane; g < uint32_t(S::kGroups) * 128; g += 32) {
                    const uint32_t a = g >> 7, kc = (g >> 3) & 15, b = g & 7;
                    const uint64_t i0 = start + uint64_t(16 * b + a) * kRowLen + 4 * kc;
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (i0 + 3 < total_samples) {
                        v = __ldg(reinterpret_cast<const float4*>(samples + i0));
                    } else {
                        if (i0 < total_samples) v.x = samples[i0];
                        if (i0 + 1 < total_samples) v.y = samples[i0 + 1];
                        if (i0 + 2 < total_samples) v.z = samples[i0 + 2];
                    }
                    *reinterpret_cast<float4*>(dst + a * kGroupBytes + (kc >> 3) * kAtomBytes + b * 128 + (((kc & 7) ^ b) << 4)) = v;
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm.full[s]);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ================================================================ MMA issuer
        uint32_t cnt[kTeams] = {};
        uint32_t j = 0;
        TileDesc nxt = load_desc(blockIdx.x);
        for (uint32_t tl = blockIdx.x; tl < n_tiles; tl += gridDim.x, ++j) {
            const TileDesc td = nxt;
            nxt = load_desc(tl + gridDim.x);
            if (td.kind != kKind) continue;
            const uint32_t s = j % kTeams;
            uint32_t ph = 0;
#pragma unroll
            for (int k = 0; k < kTeams; ++k)
                if (uint32_t(k) == s) ph = cnt[k]++ & 1u;
            mbar_wait(&sm.full[s], ph);
            mbar_wait(&sm.tempty[s], ph ^ 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (elect_one()) {
                const uint64_t ad = umma_desc(smem_u32(stage0 + s * kStageBytes), kGroupBytes), bd = umma_desc(smem_u32(tmat), kAtomBytes);
                const uint32_t alo = uint32_t(ad), ahi = uint32_t(ad >> 32), blo = uint32_t(bd), bhi = uint32_t(bd >> 32);
                const uint32_t d = tmem_base + s * kRowLen;
                uint32_t acc = 0;
                // (the start-address field holds address >> 4 in 14 bits; shared memory ends below 2^18: no carry out of it)
#pragma unroll
                for (int h = 0; h < S::kHalves; ++h) {
                    const int ksteps = (P + 63 - 64 * h) >= 64 ? 8 : (P + 63 - 64 * h + 7) / 8;  // taps reach k < P + 63
#pragma unroll
                    for (int kk = 0; kk < ksteps; ++kk) {  // K = 8 samples per instruction: column block kk / 4, 32 B further per step
                        umma_tf32_lohi(d, alo + ((h * kGroupBytes + (kk >> 2) * kAtomBytes + (kk & 3) * 32) >> 4), ahi,
                                       blo + ((h * kTBytes + (kk >> 2) * 8192 + (kk & 3) * 32) >> 4), bhi, kIdesc, acc);
                        acc = 1;
                    }
                }
                umma_commit(&sm.tfull[s]);
            }
            __syncwarp();
        }
    } else {
        // ================================================================ epilogue: energy, candidate test, block statistics
        const uint32_t team = uint32_t(warp - 2) >> 2;
        const int q = warp & 3;            // TMEM lane quarter of this warp
        const int m = 32 * q + lane;       // MMA row = TMEM lane
        const int a = m >> 3, b = m & 7;
        const int rho = 16 * b + a;        // capture row of the tile
        const uint32_t s = team;
        uint32_t cnt = 0;
        TileDesc nxt = load_desc(blockIdx.x + team * gridDim.x);
        for (uint32_t tl = blockIdx.x + team * gridDim.x; tl < n_tiles; tl += kTeams * gridDim.x) {
            const TileDesc td = nxt;
            nxt = load_desc(tl + kTeams * gridDim.x);
            if (td.kind != kKind) continue;
            const uint32_t ph = cnt++ & 1u;
            const uint64_t nwin = td.nwin;  // windows from the tile's first one on
            const float thr_c = td.thr * TMK_CAND_RATIO;
            const float kap = thr_c * thr_c * float(P);
            const uint64_t wrow = uint64_t(rho) * kRowLen;  // first window of this thread's row, from the tile start
            const int nvalid = wrow >= nwin ? 0 : (nwin - wrow < uint64_t(kRowLen) ? int(nwin - wrow) : kRowLen);
            const bool full = nwin >= uint64_t(TMK_TILE);
            mbar_wait(&sm.full[s], ph);    // the samples are visible to this thread
            const uint8_t* row = stage0 + s * kStageBytes + a * kGroupBytes + b * 128;  // the row (and, + kGroupBytes, the rows behind it)
            const uint32_t taddr = tmem_base + (uint32_t(32 * q) << 16) + s * kRowLen;
            RowStat o{__int_as_float(0x7f800000), 0.0f, 0.0f, 0u};
            uint32_t mask[2];
            float* dbg = dbg_dot != nullptr ? dbg_dot + td.start + wrow : nullptr;
            if (full) epi_row<P, true>(row, b, taddr, kap, kRowLen, &sm.tfull[s], ph, &sm.tempty[s], lane, dbg, o, mask);
            else epi_row<P, false>(row, b, taddr, kap, nvalid, &sm.tfull[s], ph, &sm.tempty[s], lane, dbg, o, mask);
            // ---- the staged samples are no longer needed
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.empty[s]);
            // ---- candidate words: the row owns words 2 (rho % 4), 2 (rho % 4) + 1 of its block of 4 rows
            const uint32_t blk = uint32_t(rho) >> 2;
            if (uint64_t(blk) * TMK_BLK < nwin) {
                uint2* wp = reinterpret_cast<uint2*>(cand_bits + (td.blk0 + blk) * TMK_BLK_WORDS + 2 * (rho & 3));
                *wp = make_uint2(mask[0], mask[1]);
            }
            // ---- block statistics: the four rows a = 4 q .. 4 q + 3 of one b ARE a block (lanes b, b + 8, b + 16, b + 24)
            float ncf = float(o.ncand);
#pragma unroll
            for (int sft = 8; sft <= 16; sft <<= 1) {
                o.emin = fminf(o.emin, __shfl_xor_sync(kFull, o.emin, sft));
                o.emax = max_nan(o.emax, __shfl_xor_sync(kFull, o.emax, sft));
                o.rmax = fmaxf(o.rmax, __shfl_xor_sync(kFull, o.rmax, sft));
                ncf += __shfl_xor_sync(kFull, ncf, sft);
            }
            if (lane < 8 && uint64_t(blk) * TMK_BLK < nwin) {  // lane b holds block 4 b + q
                // max corr^2 with the TF32 margin (see the header): 2 x 2^-10 of it, rounded up
                blk_info[td.blk0 + blk] = make_float4(o.emin, o.emax, o.rmax * (1.0f / float(P)) * 1.0015f, ncf);
                if (ncf > 0.f && cblk_list) cblk_list[atomicAdd(cblk_count, 1u)] = uint32_t(td.blk0 + blk);
            }
        }
    }
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kAccStages * kRowLen) : "memory");
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        cudaGetLastError();
    }
    return fn;
}

// the box of one tile: (32 samples = 128 B) x (8 rows b, 16 capture rows apart) x (2 column blocks of a row) x (row
// groups a), written with the 128-byte swizzle: group a lands as two swizzle atoms [columns 0..31 | 32..63] of 8 rows
bool make_map(CUtensorMap* map, const float* base, uint64_t n_samples, int groups) {
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr || getenv("TMK_SCAN_NO_TMA")) return false;
    if (n_samples < 32) return false;
    const cuuint64_t gdim[4] = {n_samples, 8, 2, cuuint64_t(groups)};
    const cuuint64_t gstr[3] = {16 * kRowLen * 4, 32 * 4, kRowLen * 4};
    const cuuint32_t box[4] = {32, 8, 2, cuuint32_t(groups)};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    const CUresult rc = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), gdim, gstr, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) {
        static bool said = false;
        if (!said) fprintf(stderr, "libtmkphy: cuTensorMapEncodeTiled failed (%d): the scan stages its tiles without TMA\n", int(rc));
        said = true;
        return false;
    }
    return true;
}

template <int LC>
int launch_tc(const float* samples, uint64_t total_samples, const TileDesc* tdesc, uint32_t n_tiles, float4* blk_info,
              uint32_t* cand_bits, uint32_t* cblk_list, unsigned int* cblk_count, float* dbg_dot, cudaStream_t st) {
    using S = Tc<LC>;
    const size_t smem = size_t(kStages) * S::kGroups * kGroupBytes + size_t(S::kHalves) * 16384 + sizeof(Smem) + 1024;
    static TmkPerDevice attr;  // (one per instantiation)
    int sms = 0, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (!attr.done(dev)) {
        cudaFuncSetAttribute(scan_tc_kernel<LC>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        attr.set(dev);
    }
    // one tensor map per window of kMapWin samples (the box coordinate is a signed 32-bit sample index); a tile is
    // loaded through the map of the window its first sample lies in, which therefore reaches one tile past the window
    TcMaps maps;
    std::memset(&maps, 0, sizeof maps);
    int n_maps = 0;
    for (uint64_t base = 0; base < total_samples && n_maps < kMaxMaps; base += kMapWin) {
        uint64_t len = total_samples - base;
        if (len > kMapWin + 16384) len = kMapWin + 16384;
        if (!make_map(&maps.m[n_maps], samples + base, len, S::kGroups)) break;
        ++n_maps;
    }
    const uint32_t grid = n_tiles < uint32_t(sms) ? n_tiles : uint32_t(sms);
    scan_tc_kernel<LC><<<grid, kThreads, smem, st>>>(maps, n_maps, samples, total_samples, tdesc, n_tiles, blk_info, cand_bits,
                                                      cblk_list, cblk_count, dbg_dot);
    return 1;
}

}  // namespace

size_t tmk_scan_tile_desc_bytes(uint32_t n_tiles) { return sizeof(TileDesc) * size_t(n_tiles ? n_tiles : 1); }

// kinds: bit 0 = some stream is 4B5B/3/2, bit 1 = Manchester/3/2, bit 2 = any other shape (generic kernel)
// tile_desc: scratch of tmk_scan_tile_desc_bytes(n_tiles) bytes (descriptors of the tiles [tile0, tile0 + n_tiles))
int tmk_launch_scan(const float* samples, uint64_t total_samples, const TmkStreamDev* streams, uint32_t n_streams,
                    const uint32_t* tile_stream, uint32_t tile0, uint32_t n_tiles, const TmkPattern* patterns,
                    uint32_t kinds, float4* blk_info, uint32_t* cand_bits, uint32_t* cblk_list, unsigned int* cblk_count,
                    cudaStream_t st, void* tile_desc, float* dbg_dot, bool build_dir) {
    if (n_tiles == 0) return 0;
    int launches = 0;
    TileDesc* td = static_cast<TileDesc*>(tile_desc);
    if (kinds & 3u) {
        tile_desc_kernel<<<(n_tiles + 127) / 128, 128, 0, st>>>(streams, n_streams, tile_stream,
                                                                build_dir && tile0 == 0 ? const_cast<uint32_t*>(tile_stream) : nullptr, patterns,
                                                                tile0, n_tiles, td);
        ++launches;
    }
    if (kinds & 1u)
        launches += launch_tc<TMK_LINE_4B5B>(samples, total_samples, td, n_tiles, blk_info, cand_bits, cblk_list, cblk_count, dbg_dot, st);
    if (kinds & 2u)
        launches += launch_tc<TMK_LINE_MANCHESTER>(samples, total_samples, td, n_tiles, blk_info, cand_bits, cblk_list, cblk_count,
                                                   dbg_dot, st);
    if (kinds & 4u) launches += tmk_launch_scan_generic(samples, streams, tile_stream, tile0, n_tiles, patterns, blk_info, cand_bits,
                                                            cblk_list, cblk_count, st);
    return launches;
}
#endif  // !TMK_SCAN_ARM_A
