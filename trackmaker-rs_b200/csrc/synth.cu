// synth.cu — synthetic capture generator for the measurement harness: every stream
// is AWGN plus randomly placed src/phy frames (random payloads, 10 % ACKs, 10 % foreign dst,
// gaps U[48,4800)), generated on the GPU from counter-based hashes of (seed, stream, index).
// The frames are produced exactly as PhyEncoder would (encoder.rs:43-63): header + CRC-8 +
// line code behind the preamble, scaled by `amplitude` and added onto the noise.
#include "tmk_internal.cuh"

namespace {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__constant__ uint8_t c_enc5s[16] = {0x1e, 0x09, 0x14, 0x15, 0x0a, 0x0b, 0x0e, 0x0f,
                                    0x12, 0x13, 0x16, 0x17, 0x1a, 0x1b, 0x1c, 0x1d};

// 4 samples per thread: two Box-Muller pairs
__global__ void noise_kernel(uint64_t seed, uint64_t total, float sigma, float* __restrict__ out) {
    const uint64_t i4 = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
    if (i4 >= total) return;
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
                            uint64_t n_per_stream, uint32_t* __restrict__ counts, const uint64_t* __restrict__ base,
                            tmk_frame_rec* __restrict__ frames, uint64_t frame_cap) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
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
        if (pos + flen > n_per_stream) break;
        if (frames) {
            const uint64_t slot = base[s] + k;
            if (slot < frame_cap) {
                tmk_frame_rec t;
                t.stream = s;
                t.status = TMK_FRAME_OK;
                t.preamble_start = pos;
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
        pos += flen + 48 + next() % (4800 - 48);
    }
    if (!frames) counts[s] = k;
}

__device__ __forceinline__ uint8_t payload_byte(uint64_t seed, uint32_t stream, uint64_t pos, uint32_t i) {
    return uint8_t(mix64(seed ^ (uint64_t(stream) << 40) ^ (pos << 8) ^ 0xABCDull ^ (uint64_t(i) << 52)) >> 17);
}

// One CTA per planned frame: build the frame bytes, then add amplitude * level onto the capture.
__global__ void __launch_bounds__(128)
frames_kernel(const TmkPattern* __restrict__ pat, uint64_t seed, uint64_t n_per_stream, float amplitude,
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
    float* __restrict__ o = samples + uint64_t(fr.stream) * n_per_stream + fr.preamble_start;
    const uint32_t total = P + nlev * spl;
    for (uint32_t i = threadIdx.x; i < total; i += blockDim.x) {
        const float level = i < P ? pat->levels[i / spl] : (lvl[(i - P) / spl] ? -1.0f : 1.0f);
        o[i] += amplitude * level;
    }
}

}  // namespace

int tmk_launch_synth_noise(uint64_t seed, uint64_t total, float sigma, float* samples, cudaStream_t st) {
    if (total == 0) return 0;
    const uint64_t threads = (total + 3) / 4;
    noise_kernel<<<uint32_t((threads + 255) / 256), 256, 0, st>>>(seed, total, sigma, samples);
    return 1;
}

int tmk_launch_synth_plan(const TmkPattern* pattern, int32_t local_addr, uint64_t seed, uint32_t n_streams,
                          uint64_t n_per_stream, uint32_t* counts, const uint64_t* base, tmk_frame_rec* frames,
                          uint64_t frame_cap, cudaStream_t st) {
    if (n_streams == 0) return 0;
    plan_kernel<<<(n_streams + 127) / 128, 128, 0, st>>>(pattern, local_addr, seed, n_streams, n_per_stream, counts,
                                                        base, frames, frame_cap);
    return 1;
}

int tmk_launch_synth_frames(const TmkPattern* pattern, uint64_t seed, uint64_t n_per_stream, float amplitude,
                            tmk_frame_rec* frames, uint64_t n_frames, float* samples, cudaStream_t st) {
    if (n_frames == 0) return 0;
    frames_kernel<<<uint32_t(n_frames), 128, 0, st>>>(pattern, seed, n_per_stream, amplitude, frames, samples);
    return 1;
}
