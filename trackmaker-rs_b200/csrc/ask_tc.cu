// ask_tc.cu — the chirp filter of the examples/ask.rs receiver on the 5th-generation tensor cores.
//
// Replaces, as a FILTER, the per-sample work of the RX loop (examples/ask.rs:693-714: a 440-tap sliding dot product / 200
// against the chirp, compared with twice the EMA power and 0.05): it marks every 32-sample word in which
// `sync_power > 2 power && sync_power > 0.05` could hold; the walk (ask.cu) evaluates marked words exactly, in the
// reference's f32 order, so reduced precision here cannot change a result. The FFT fast correlation (ask.cu,
// ask_scan_fft_kernel: 22 ms at C2, shared-memory bound) stays as the A/B arm behind TMK_ASK_FFT=1.
//
//   The capture is read as rows of 64 samples; a tile is 128 rows (8192 window starts). With n the window start
//   (position i = n + 439 is the sample the reference has just shifted in):
//   D[row r][offset n] = sum_t x[64 r + n + t] h[t] = sum_{h=0..7} X_h (128 x 64) . T_h^T (64 x 64),
//   X_h[r][k] = x[64 (r + h) + k], T_h[n][k] = chirp[64 h + k - n]  (constant, 8 x 16 KB of shared memory).
//   TMA      one 4-D box per tile (23 row groups, 128-byte swizzle) lands the rows in tcgen05's canonical K-major layout
//            with MMA row i = 8 a + b holding capture row 16 b + a, so X_h is the same buffer with the operand address
//            moved by h row groups (the layout of scan_tc.cu).
//   UMMA     tcgen05.mma kind::tf32, M 128, N 64, K 8: 63 instructions per tile, one issuing thread, TMEM accumulators.
//   epilogue two teams of four warps: a thread owns a row (64 consecutive positions): the EMA power runs from the exact
//            checkpoint of its word (ask_power_kernel; unknown -> 0, which only adds candidates) over the staged samples,
//            the margin comes from the energy of the 8 rows under its windows, marked words are OR-ed into the flags.
//   TF32 keeps 10 mantissa bits of both operands: |D_tf32 - D| <= 2^-9 (1 + 2^-10) ||x_window|| ||chirp||, i.e.
//   <= 1.46e-4 sqrt(E) in units of sync_power (||chirp|| = 14.83, / 200), E <= the energy of the 8 rows; the walk's own
//   sequential f32 sum differs from the real value by at most 440 u of the same bound (2e-6 sqrt(E)). Margin: 1.6e-4 sqrt(E).
// Bound: HBM (4 B per sample read once; 2.0 TFLOP of TF32 at C2 is ~2 ms of the tensor pipe).
#include <cuda.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "engine.hpp"

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kPre = TMK_ASK_PREAMBLE;         // 440
constexpr int kRowLen = 64, kRows = 128;
constexpr int kTile = kRows * kRowLen;         // 8192 window starts per tile
constexpr int kHalves = 8;                     // taps reach sample offset 63 + 439 = 502 < 8 x 64
constexpr int kGroups = 16 + kHalves - 1;      // row groups a per stage
constexpr int kTeams = 2, kStages = 4, kAccStages = 4;  // tile j of a CTA: team j % 2, shared-memory stage and TMEM stage j % 4
constexpr int kSlices = (kPre + 63 + 7) / 8;            // K slices of 8 taps that reach a tap: 63
constexpr int kWBlocks = kSlices + 7;                   // blocks of the Toeplitz strip (see the kernel): 70 x 256 B
constexpr int kEpiWarps = 4 * kTeams;
constexpr int kThreads = 32 * (2 + kEpiWarps);
constexpr uint32_t kGroupBytes = 2048, kAtomBytes = 1024, kWBytes = kWBlocks * 256;
constexpr uint32_t kStageBytes = kGroups * kGroupBytes;
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(kRowLen >> 3) << 17) | (uint32_t(kRows >> 4) << 24);
constexpr float kC = 0.984375f, kInv64 = 0.015625f;
constexpr int kMaxMaps = 20;
constexpr uint64_t kMapWin = (1ull << 31) - (1ull << 16);
struct TcMaps {
    CUtensorMap m[kMaxMaps];
};

struct alignas(16) TileDescA {
    uint64_t start;  // first sample of the tile in the batch buffer (= its first window start)
    uint64_t nwin;   // valid window starts from the tile's first one on (n + 439 < N)
    uint64_t pos0;   // 32 x (global word index) + sample in the word, of the position of the tile's first window: 32 word_base + 8192 t + 439
    uint64_t pad_;
};
static_assert(sizeof(TileDescA) == 32, "tile descriptor");

// tile -> descriptor: tc_base[s] = first tile of stream s (prefix over the streams), binary search per tile
__global__ void ask_tile_desc_kernel(const uint64_t* __restrict__ s_off, const uint64_t* __restrict__ s_n, const uint64_t* __restrict__ s_word,
                                     const uint32_t* __restrict__ tc_base, uint32_t n_streams, uint32_t n_tiles, TileDescA* __restrict__ out) {
    const uint32_t tl = blockIdx.x * blockDim.x + threadIdx.x;
    if (tl >= n_tiles) return;
    uint32_t lo = 0, hi = n_streams - 1;
    while (lo < hi) {
        const uint32_t mid = (lo + hi + 1) >> 1;
        if (tc_base[mid] <= tl) lo = mid; else hi = mid - 1;
    }
    const uint64_t w0 = uint64_t(tl - tc_base[lo]) * kTile;
    TileDescA d;
    d.start = s_off[lo] + w0;
    d.nwin = s_n[lo] - (kPre - 1) - w0;  // (a stream has tiles only while this is positive)
    d.pos0 = s_word[lo] * 32 + w0 + (kPre - 1);
    d.pad_ = 0;
    out[tl] = d;
}

struct Smem {
    uint64_t full[kStages], empty[kStages], tfull[kAccStages], tempty[kAccStages];
    float rowsum[kTeams][kRows + 8];
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
// K-major, 128-byte swizzle operand descriptor (see scan_tc.cu: umma_desc)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t sbo) {
    return uint64_t((saddr >> 4) & 0x3fffu) | (1ull << 16) | (uint64_t((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46) | (2ull << 61);
}
// K-major, no swizzle: core matrices of 8 rows x 16 B; LBO = bytes between the two core matrices of an instruction's K = 8,
// SBO = bytes between 8-row groups
__device__ __forceinline__ uint64_t umma_desc_ns(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return uint64_t((saddr >> 4) & 0x3fffu) | (uint64_t((lbo >> 4) & 0x3fffu) << 16) | (uint64_t((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// the same with the descriptors as (low word, high word): per instruction only the low words change, by a constant — one add
// each, where the 64-bit form recomputed shift, mask and merge (13 SASS instructions and ~49 cycles per MMA in the one
// issuing thread: the filter was bound by the ISSUE of its 63 MMAs per tile, not by the tensor pipe)
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

__global__ void __launch_bounds__(kThreads, 1)
ask_tc_kernel(const __grid_constant__ TcMaps maps, int n_maps, const float* __restrict__ samples, uint64_t total_samples,
              const TileDescA* __restrict__ tdesc, uint32_t n_tiles, const float* __restrict__ chirp, const float* __restrict__ pchk,
              uint32_t* __restrict__ wflag) {
    extern __shared__ uint8_t smem_dyn[];
    uint8_t* smem_raw = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    uint8_t* stage0 = smem_raw;
    uint8_t* tmat = smem_raw + kStages * kStageBytes;
    Smem& sm = *reinterpret_cast<Smem*>(tmat + ((kWBytes + 127u) & ~127u));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&sm.full[i], 1);
            mbar_init(&sm.empty[i], 4);
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
    // The Toeplitz operand of K slice s (taps 8 s .. 8 s + 7): B_s[n][kk] = chirp[8 s + kk - n], n = 8 u + v
    //   = chirp[8 (s - u) + kk - v]: a function of w = s - u and (v, kk) only. So ONE strip of 8 x 8 blocks
    //   W_w[v][kk] = chirp[8 w + kk - v], w = 62 .. -7 (stored in that order: block j holds w = 62 - j), serves every slice:
    //   B_s is the eight consecutive blocks j = 62 - s + u, u = 0 .. 7 — the operand descriptor of slice s points at block 62 - s
    //   with a stride of one block between the 8-row groups. 70 blocks x 256 B = 17.5 KB instead of 8 x 16 KB of explicit
    //   T_h, which leaves room for FOUR stages of samples (the kernel was bound by its two-stage pipeline: 4.0 ms at C2
    //   with 2.5 ms of MMA). Layout of a block: tcgen05's K-major layout without swizzle, two core matrices (taps 0-3, 4-7)
    //   of 8 rows x 16 B.
    for (uint32_t i = threadIdx.x; i < uint32_t(kWBlocks) * 64; i += kThreads) {
        const uint32_t j = i >> 6, v = (i >> 3) & 7, kk = i & 7;
        const int t = 8 * (kSlices - 1 - int(j)) + int(kk) - int(v);
        const float val = (t >= 0 && t < kPre) ? __ldg(chirp + t) : 0.0f;
        *reinterpret_cast<float*>(tmat + j * 256 + (kk >> 2) * 128 + v * 16 + (kk & 3) * 4) = val;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = sm.tmem_base;

    auto load_desc = [&](uint32_t tl) -> TileDescA {
        TileDescA d;
        d.start = d.nwin = d.pos0 = d.pad_ = 0;
        if (tl < n_tiles) {
            const uint4* p = reinterpret_cast<const uint4*>(tdesc + tl);
            const uint4 lo = __ldg(p), hi = __ldg(p + 1);
            d.start = uint64_t(lo.x) | (uint64_t(lo.y) << 32);
            d.nwin = uint64_t(lo.z) | (uint64_t(lo.w) << 32);
            d.pos0 = uint64_t(hi.x) | (uint64_t(hi.y) << 32);
        }
        return d;
    };

    // iteration j of a CTA is tile blockIdx.x + j gridDim.x; team j % kTeams owns it with stage and TMEM stage j % kTeams
    if (warp == 0) {
        // ================================================================ producer
        uint32_t j = 0;
        TileDescA nxt = load_desc(blockIdx.x);  // (the descriptor of the next tile is in flight while this one is staged)
        for (uint32_t tl = blockIdx.x; tl < n_tiles; tl += gridDim.x, ++j) {
            const TileDescA td = nxt;
            nxt = load_desc(tl + gridDim.x);
            const uint32_t s = j % kStages, ph = (j / kStages) & 1u;
            mbar_wait(&sm.empty[s], ph ^ 1u);
            uint8_t* dst = stage0 + s * kStageBytes;
            const uint64_t start = td.start;
            const uint64_t span = uint64_t(16 * 7 + kGroups) * kRowLen;
            const uint32_t mi = uint32_t(start / kMapWin);
            const bool tma = mi < uint32_t(n_maps) && start + span <= total_samples;
            if (tma) {
                if (lane == 0) {
                    mbar_expect_tx(&sm.full[s], kStageBytes);
                    const int32_t c0 = int32_t(start - uint64_t(mi) * kMapWin);
                    asm volatile(
                        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(&maps.m[mi])), "r"(smem_u32(&sm.full[s])), "r"(c0), "r"(0),
                        "r"(0), "r"(0) : "memory");
                }
            } else {  // the last tiles of the buffer and unsupported maps: the warp fills the same layout itself
                for (uint32_t g = lane; g < uint32_t(kGroups) * 128; g += 32) {
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
        uint32_t j = 0;
        for (uint32_t tl = blockIdx.x; tl < n_tiles; tl += gridDim.x, ++j) {
            const uint32_t s = j % kStages, ph = (j / kStages) & 1u;  // (kAccStages == kStages: the same index and phase)
            mbar_wait(&sm.full[s], ph);
            mbar_wait(&sm.tempty[s], ph ^ 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (elect_one()) {
                const uint64_t ad = umma_desc(smem_u32(stage0 + s * kStageBytes), kGroupBytes), bd = umma_desc_ns(smem_u32(tmat), 128, 256);
                const uint32_t alo = uint32_t(ad), ahi = uint32_t(ad >> 32), blo = uint32_t(bd), bhi = uint32_t(bd >> 32);
                const uint32_t d = tmem_base + s * kRowLen;
                uint32_t acc = 0;
                // (the start-address field holds address >> 4 in 14 bits; shared memory ends below 2^18, so adding a byte
                //  offset >> 4 to the low word never carries out of the field)
#pragma unroll
                for (int h = 0; h < kHalves; ++h) {
                    const int left = kPre + 63 - 64 * h;  // taps reach k < left in this half
                    const int ksteps = left >= 64 ? 8 : (left + 7) / 8;
#pragma unroll
                    for (int kk = 0; kk < ksteps; ++kk) {
                        umma_tf32_lohi(d, alo + ((h * kGroupBytes + (kk >> 2) * kAtomBytes + (kk & 3) * 32) >> 4), ahi,
                                       blo + ((uint32_t(kSlices - 1 - (8 * h + kk)) * 256u) >> 4), bhi, kIdesc, acc);
                        acc = 1;
                    }
                }
                umma_commit(&sm.tfull[s]);
            }
            __syncwarp();
        }
    } else {
        // ================================================================ epilogue
        const uint32_t team = uint32_t(warp - 2) >> 2;
        const int q = warp & 3;        // TMEM lane quarter of this warp (warp % 4)
        const int m = 32 * q + lane;   // MMA row = TMEM lane
        const int a = m >> 3, b = m & 7;
        const int rho = 16 * b + a;    // capture row of the tile
        uint32_t cnt = 0;
        float* rs = sm.rowsum[team];
        // The descriptor of a tile and, behind it, the power checkpoint of the row are two dependent trips to L2 / HBM: both are
        // fetched one and two tiles ahead (47 % of the kernel's samples sat on the first use of the checkpoint before)
        const uint64_t wrow = uint64_t(rho) * kRowLen;
        auto valid_of = [&](const TileDescA& d) -> int {
            return wrow >= d.nwin ? 0 : (d.nwin - wrow < uint64_t(kRowLen) ? int(d.nwin - wrow) : kRowLen);
        };
        // the exact power before the first sample of the word that holds the row's first position
        auto chk_of = [&](const TileDescA& d) -> float { return valid_of(d) > 0 ? __ldg(pchk + ((d.pos0 + wrow) >> 5)) : 0.0f; };
        const uint32_t step = kTeams * gridDim.x;
        uint32_t tl = blockIdx.x + team * gridDim.x;
        TileDescA td_cur = load_desc(tl), td_nxt = load_desc(tl + step);
        float p_cur = chk_of(td_cur);
        for (; tl < n_tiles; tl += step) {
            const TileDescA td = td_cur;
            float p = p_cur;
            td_cur = td_nxt;
            p_cur = chk_of(td_cur);
            td_nxt = load_desc(tl + 2 * step);
            const uint32_t j = team + kTeams * cnt++;  // this CTA's tile counter
            const uint32_t s = j % kStages, ph = (j / kStages) & 1u;
            const int nvalid = valid_of(td);
            const uint64_t bit0 = td.pos0 + wrow;  // 32 x global word + sample in the word, of the row's first position (bit0 % 32 == 23)
            mbar_wait(&sm.full[s], ph);
            const uint8_t* stage = stage0 + s * kStageBytes;
            const uint8_t* row = stage + a * kGroupBytes + b * 128;
            auto chunk = [&](int c) -> float4 {  // samples 4 c .. 4 c + 3 counted from the row start (rows behind it follow)
                return *reinterpret_cast<const float4*>(row + (c >> 4) * kGroupBytes + ((c >> 3) & 1) * kAtomBytes + (((c & 7) ^ b) << 4));
            };
            // ---- energy of every row of the stage (own row; the 7 rows behind the tile by the threads rho < 7)
            {
                float e = 0.0f;
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const float4 v = chunk(c);
                    e = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, e))));
                }
                rs[rho] = e;
                if (rho < kHalves - 1) {  // capture row 128 + rho = 16 * 7 + (16 + rho): group 16 + rho, b = 7
                    const uint8_t* xr = stage + (16 + rho) * kGroupBytes + 7 * 128;
                    float e2 = 0.0f;
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const float4 v = *reinterpret_cast<const float4*>(xr + ((c >> 3) & 1) * kAtomBytes + (((c & 7) ^ 7) << 4));
                        e2 = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, e2))));
                    }
                    rs[kRows + rho] = e2;
                }
            }
            asm volatile("bar.sync %0, 128;" ::"r"(1 + int(team)) : "memory");  // the team's four warps
            float e8 = 0.0f;
#pragma unroll
            for (int k = 0; k < kHalves; ++k) e8 += rs[rho + k];
            // (NaN / inf samples: the margin and every dot product of the row are NaN -> all its words are marked)
            const float margin = 1.6e-4f * sqrtf(e8) + 1e-7f;
            if (p < 0.0f) p = 0.0f;  // unknown checkpoint: a chain from 0 is a lower bound of the power, which only adds candidates
            // ---- warm-up of the EMA: the 23 samples between the word boundary and the row's first position
#pragma unroll
            for (int sidx = kPre - 1 - 23; sidx < kPre - 1; ++sidx) {
                const float4 v = chunk(sidx >> 2);
                const float x = (sidx & 3) == 0 ? v.x : (sidx & 3) == 1 ? v.y : (sidx & 3) == 2 ? v.z : v.w;
                p = fmaf(p, kC, x * x * kInv64);
            }
            mbar_wait(&sm.tfull[s], ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + (uint32_t(32 * q) << 16) + s * kRowLen;
            uint32_t mask[2];
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                float D[32];
                tmem_ld32(taddr + 32 * half, D);
                if (half == 1) {
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm.tempty[s]);
                }
                uint32_t mk = 0;
#pragma unroll
                for (int jj = 0; jj < 32; ++jj) {
                    const int sidx = kPre - 1 + 32 * half + jj;  // the sample shifted in at this position
                    const float4 v = chunk(sidx >> 2);
                    const float x = (sidx & 3) == 0 ? v.x : (sidx & 3) == 1 ? v.y : (sidx & 3) == 2 ? v.z : v.w;
                    p = fmaf(p, kC, x * x * kInv64);
                    const float sp = D[jj] * 0.005f + margin;
                    const bool cand = !(sp <= 0.05f) && !(sp <= 2.0f * p * (1.0f - 1e-4f));  // NaN -> candidate
                    if (cand && 32 * half + jj < nvalid) mk |= 1u << jj;
                }
                mask[half] = mk;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.empty[s]);
            // ---- the flags hold ONE BIT PER 32-SAMPLE WORD; the row's 64 positions start at sample 23 of word bit0 / 32
            //      and reach into the two words behind it
            if (mask[0] | mask[1]) {
                const uint64_t m64 = uint64_t(mask[0]) | (uint64_t(mask[1]) << 32);
                const uint32_t sh = 32u - uint32_t(bit0 & 31);  // positions of the row inside its first word (9)
                const uint64_t w0 = bit0 >> 5;
                const bool f0 = (m64 & ((1ull << sh) - 1ull)) != 0, f1 = ((m64 >> sh) & 0xffffffffull) != 0, f2 = (m64 >> (sh + 32)) != 0;
                if (f0) atomicOr(wflag + (w0 >> 5), 1u << (w0 & 31));
                if (f1) atomicOr(wflag + ((w0 + 1) >> 5), 1u << ((w0 + 1) & 31));
                if (f2) atomicOr(wflag + ((w0 + 2) >> 5), 1u << ((w0 + 2) & 31));
            }
            asm volatile("bar.sync %0, 128;" ::"r"(1 + int(team)) : "memory");  // rowsum is rewritten by the team's next tile
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
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        cudaGetLastError();
    }
    return fn;
}
bool make_map(CUtensorMap* map, const float* base, uint64_t n_samples) {
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr || getenv("TMK_SCAN_NO_TMA") || n_samples < 32) return false;
    const cuuint64_t gdim[4] = {n_samples, 8, 2, cuuint64_t(kGroups)};
    const cuuint64_t gstr[3] = {16 * kRowLen * 4, 32 * 4, kRowLen * 4};
    const cuuint32_t box[4] = {32, 8, 2, cuuint32_t(kGroups)};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

// Tile descriptors of the whole batch (on e->stream), then one launch per range of streams. s_off / s_n / s_word: per-stream
// sample offset, length and first flag word (host arrays); tile_base_out[s] = first tile of stream s (n_streams + 1 entries).
int tmk_ask_tc_prepare(tmk_engine* e, const uint64_t* s_off, const uint64_t* s_n, const uint64_t* s_word, uint32_t n_streams,
                       uint32_t* tile_base_out) {
    uint64_t tiles = 0;
    for (uint32_t s = 0; s < n_streams; ++s) {
        tile_base_out[s] = uint32_t(tiles);
        if (s_n[s] >= uint64_t(kPre)) tiles += (s_n[s] - (kPre - 1) + kTile - 1) / kTile;
        if (tiles > 0x7fffffffull) return tmk_fail(e, TMK_E_INVALID, "batch too large (tile count)");
    }
    tile_base_out[n_streams] = uint32_t(tiles);
    if (tiles == 0) return TMK_OK;
    const size_t arr = sizeof(uint64_t) * n_streams;
    CK(e->d_ask_tc.reserve(3 * arr + sizeof(uint32_t) * (n_streams + 1) + 64 + sizeof(TileDescA) * tiles));
    char* d = static_cast<char*>(e->d_ask_tc.p);
    CK(cudaMemcpyAsync(d, s_off, arr, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(d + arr, s_n, arr, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(d + 2 * arr, s_word, arr, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(d + 3 * arr, tile_base_out, sizeof(uint32_t) * (n_streams + 1), cudaMemcpyHostToDevice, e->stream));
    TileDescA* td = reinterpret_cast<TileDescA*>(d + (3 * arr + sizeof(uint32_t) * (n_streams + 1) + 31) / 32 * 32);
    ask_tile_desc_kernel<<<uint32_t((tiles + 255) / 256), 256, 0, e->stream>>>(
        reinterpret_cast<const uint64_t*>(d), reinterpret_cast<const uint64_t*>(d + arr), reinterpret_cast<const uint64_t*>(d + 2 * arr),
        reinterpret_cast<const uint32_t*>(d + 3 * arr), n_streams, uint32_t(tiles), td);
    e->launches += 1;
    CK(cudaGetLastError());
    return TMK_OK;
}

// Marks the candidate words of the tiles [tile0, tile1) in wflag (zeroed by the caller), on `st`. pchk: the exact power
// checkpoints (ask_power_kernel) indexed by global word.
int tmk_ask_tc_filter(tmk_engine* e, const float* dx, uint64_t total_samples, uint32_t n_streams, uint32_t tile0, uint32_t tile1,
                      const float* chirp, const float* pchk, uint32_t* wflag, cudaStream_t st) {
    if (tile1 <= tile0) return TMK_OK;
    const size_t arr = sizeof(uint64_t) * n_streams;
    const TileDescA* td = reinterpret_cast<const TileDescA*>(static_cast<char*>(e->d_ask_tc.p) +
                                                             (3 * arr + sizeof(uint32_t) * (n_streams + 1) + 31) / 32 * 32);
    const size_t smem = size_t(kStages) * kStageBytes + ((kWBytes + 127u) & ~127u) + sizeof(Smem) + 1024;
    static TmkPerDevice attr;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device));
    if (!attr.done(e->device)) {
        CK(cudaFuncSetAttribute(ask_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        attr.set(e->device);
    }
    TcMaps maps;
    std::memset(&maps, 0, sizeof maps);
    int n_maps = 0;
    for (uint64_t b0 = 0; b0 < total_samples && n_maps < kMaxMaps; b0 += kMapWin) {
        uint64_t len = total_samples - b0;
        if (len > kMapWin + 16384) len = kMapWin + 16384;
        if (!make_map(&maps.m[n_maps], dx + b0, len)) break;
        ++n_maps;
    }
    const uint32_t tiles = tile1 - tile0;
    const uint32_t grid = uint32_t(std::min<uint64_t>(tiles, uint64_t(sms)));
    ask_tc_kernel<<<grid, kThreads, smem, st>>>(maps, n_maps, dx, total_samples, td + tile0, tiles, chirp, pchk, wflag);
    e->launches += 1;
    CK(cudaGetLastError());
    return TMK_OK;
}
