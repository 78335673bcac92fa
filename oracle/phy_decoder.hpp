// ORACLE — TEST INFRASTRUCTURE ONLY (see phy_core.hpp header).
// CPU restatement of PhyEncoder (src/phy/encoder.rs:5-97) and of the streaming PhyDecoder
// state machine (src/phy/decoder.rs:11-418). Every verdict is also reported through a
// side-channel record (tmk_frame_rec) with ABSOLUTE sample positions so the CUDA engine's
// verdicts can be compared one by one.
#pragma once
#include <immintrin.h>

#include "../include/tmkphy.h"
#include "phy_core.hpp"

namespace orc {

struct Encoder {  // encoder.rs:5-97
    std::unique_ptr<LineCode> lc;
    std::vector<float> preamble;
    Encoder(size_t spl, size_t pbytes, int kind) : lc(make_line_code(kind, spl)) {
        preamble = lc->generate_preamble(pbytes);
    }
    std::vector<float> encode_frame(const Frame& f) const {  // :43-63
        std::vector<uint8_t> bytes = f.to_bytes();
        std::vector<float> body = lc->encode(bytes_to_bits(bytes.data(), bytes.size()));
        std::vector<float> out(preamble);
        out.insert(out.end(), body.begin(), body.end());
        return out;
    }
    std::vector<float> encode_frames(const std::vector<Frame>& fs, size_t gap) const {  // :70-92
        std::vector<float> out;
        for (size_t i = 0; i < fs.size(); ++i) {
            std::vector<float> s = encode_frame(fs[i]);
            out.insert(out.end(), s.begin(), s.end());
            if (i + 1 < fs.size()) out.insert(out.end(), gap, 0.0f);  // gap between, not after
        }
        return out;
    }
};

// decoder.rs:381-417 — the AVX order is the reference order on every AVX host:
// 8 lanes of mul-then-add, lane l + lane l+4, the 4 partials summed left to right from 0,
// then the len%8 tail left to right.
inline float dot_avx_order(const float* w, const float* p, size_t len) {
    __m256 acc = _mm256_setzero_ps();
    size_t i = 0;
    for (; i + 8 <= len; i += 8)
        acc = _mm256_add_ps(acc, _mm256_mul_ps(_mm256_loadu_ps(w + i), _mm256_loadu_ps(p + i)));
    __m128 q = _mm_add_ps(_mm256_castps256_ps128(acc), _mm256_extractf128_ps(acc, 1));
    float a[4];
    _mm_storeu_ps(a, q);
    float s = 0.0f;
    for (int k = 0; k < 4; ++k) s = s + a[k];
    for (; i < len; ++i) s = s + w[i] * p[i];
    return s;
}

inline float seq_energy(const float* x, size_t n) {  // iter().map(|x| x*x).sum::<f32>()
    float e = 0.0f;
    for (size_t i = 0; i < n; ++i) e = e + x[i] * x[i];
    return e;
}

struct Decoder {  // decoder.rs:16-33
    std::unique_ptr<LineCode> lc;
    std::vector<float> preamble;
    float threshold, preamble_energy;
    std::vector<float> buf;
    size_t off = 0;  // buffer_offset
    bool decoding = false;
    size_t frame_start = 0;  // DecoderState::Decoding(usize)
    size_t max_frame_bytes;
    int local_addr;
    uint64_t drained = 0;  // samples removed from the front of buf so far (for absolute positions)
    // side channel
    uint32_t stream_id = 0;
    float det_corr = 0.0f;
    uint64_t exact_calls = 0;
    std::vector<tmk_frame_rec>* recs = nullptr;
    std::vector<uint8_t>* payload = nullptr;
    std::vector<Frame> frames;  // decoded_frames of the last process_samples call

    explicit Decoder(const tmk_phy_config& c)
        : lc(make_line_code(c.line_code, c.spl)),
          threshold(c.threshold),
          max_frame_bytes(c.max_frame_bytes),
          local_addr(c.local_addr) {
        preamble = lc->generate_preamble(c.preamble_bytes);
        preamble_energy = std::sqrt(seq_energy(preamble.data(), preamble.size()));  // :46-50
    }

    void reset() {  // :114-119
        buf.clear();
        off = 0;
        decoding = false;
    }

    // decoder.rs:68-112
    const std::vector<Frame>& process_samples(const float* s, size_t n) {
        frames.clear();
        buf.insert(buf.end(), s, s + n);
        for (;;) {
            long step = decoding ? decode_frame() : search();
            if (step < 0) break;  // None: need more data
            off += size_t(step);
        }
        if (off > 0) {  // :90-109 keep P-1 samples of overlap
            size_t keep = preamble.empty() ? 0 : preamble.size() - 1;
            size_t drain_end = off > keep ? off - keep : 0;
            if (drain_end > 0) {
                buf.erase(buf.begin(), buf.begin() + drain_end);
                off -= drain_end;
                drained += drain_end;
                if (decoding) frame_start = frame_start > drain_end ? frame_start - drain_end : 0;
            }
        }
        return frames;
    }

    // decoder.rs:123-238. Returns samples consumed, or -1 for None.
    long search() {
        size_t P = preamble.size();
        if (buf.size() - off < P) return -1;
        const float* area = buf.data() + off;
        size_t area_len = buf.size() - off;
        size_t windows = area_len - P + 1;
        float energy = seq_energy(area, P);  // re-seeded on every entry, :133-136
        for (size_t i = 0; i < windows; ++i) {
            float corr;
            if (energy < 1e-6f) {
                corr = 0.0f;
            } else {
                float dot = dot_avx_order(area + i, preamble.data(), P);
                corr = dot / (std::sqrt(energy) * preamble_energy);
            }
            if (corr >= threshold) {
                // sync-word refinement, :157-207
                size_t sync_len = lc->samples_for_bits(8);
                const float* sync = preamble.data() + (P - sync_len);
                float sync_energy = std::sqrt(seq_energy(sync, sync_len));
                size_t margin = lc->samples_for_bits(1);
                size_t expect = i + P - sync_len;
                size_t lo = expect > margin ? expect - margin : 0;
                size_t hi = std::min(expect + margin, area_len - sync_len);
                float best = -1.0f;
                size_t best_off = expect;
                for (size_t j = lo; j <= hi; ++j) {
                    float dot = 0.0f, we = 0.0f;
                    for (size_t t = 0; t < sync_len; ++t) {
                        dot = dot + area[j + t] * sync[t];
                        we = we + area[j + t] * area[j + t];
                    }
                    float c = (we > 1e-6f && sync_energy > 1e-6f) ? dot / (std::sqrt(we) * sync_energy) : 0.0f;
                    if (c > best) {
                        best = c;
                        best_off = j;
                    }
                }
                frame_start = off + best_off + sync_len;
                decoding = true;
                det_corr = corr;
                return long(i);  // buffer_offset now points at the preamble start
            }
            if (i + 1 < windows) {  // rolling update with clamp, :224-233
                float leaving = area[i], entering = area[i + P];
                energy = energy - leaving * leaving + entering * entering;
                if (energy < 0.0f) energy = 0.0f;
            }
        }
        return long(windows);
    }

    void emit(uint32_t status, size_t consumed, bool have_hdr, const Header& h, const uint8_t* data,
              size_t data_len) {
        if (!recs) return;
        tmk_frame_rec r;
        std::memset(&r, 0, sizeof r);
        r.stream = stream_id;
        r.status = status;
        r.preamble_start = drained + off;
        r.frame_start = drained + frame_start;
        r.corr = det_corr;
        r.consumed = uint32_t(consumed);
        r.hdr_valid = have_hdr ? 1 : 0;
        if (have_hdr) {
            r.len = uint16_t(h.len);
            r.crc = h.crc;
            r.type = h.type;
            r.seq = h.seq;
            r.src = h.src;
            r.dst = h.dst;
        }
        r.payload_off = payload ? payload->size() : 0;
        if (payload && data) payload->insert(payload->end(), data, data + data_len);
        recs->push_back(r);
    }

    // decoder.rs:242-356
    long decode_frame() {
        size_t P = preamble.size();
        size_t hs = lc->samples_for_bits(8 * kHeaderBytes);
        if (buf.size() < frame_start + hs) return -1;
        std::vector<uint8_t> hbits = lc->decode(buf.data() + frame_start, hs);
        std::vector<uint8_t> hbytes = bits_to_bytes(hbits.data(), hbits.size());
        Header h{};
        bool have7 = false;
        if (!parse_header(hbytes, &h, &have7)) {  // :266-273 consumes header_samples
            emit(TMK_FRAME_HDR_PARSE, hs, have7, h, nullptr, 0);
            decoding = false;
            return long(hs);
        }
        if ((h.type == TMK_FRAME_TYPE_DATA && h.len == 0) || h.len > max_frame_bytes) {  // :277-286
            emit(TMK_FRAME_BAD_LEN, 1, true, h, nullptr, 0);
            decoding = false;
            return 1;
        }
        size_t total_bits = (kHeaderBytes + h.len) * 8;
        size_t total_samples = lc->samples_for_bits(total_bits);
        if (buf.size() < frame_start + total_samples) return -1;
        std::vector<uint8_t> bits = lc->decode(buf.data() + frame_start, total_samples);
        size_t consumed = P + lc->samples_for_bits(bits.size());  // :306-309
        decoding = false;
        if (bits.size() < total_bits) {  // :311-320
            emit(TMK_FRAME_SHORT_DECODE, consumed, true, h, nullptr, 0);
            return long(consumed);
        }
        std::vector<uint8_t> bytes = bits_to_bytes(bits.data(), bits.size());
        const uint8_t* data = bytes.data() + kHeaderBytes;
        if (local_addr >= 0 && h.dst != local_addr) {  // :322-329
            emit(TMK_FRAME_NOT_FOR_US, consumed, true, h, data, h.len);
            return long(consumed);
        }
        // Frame::from_bits -> from_bytes, frame.rs:147-183: header re-parsed from the same
        // bytes (identical), CRC-8 over the payload only.
        if (crc8(data, h.len) != h.crc) {
            emit(TMK_FRAME_CRC_FAIL, consumed, true, h, data, h.len);
            return long(consumed);
        }
        Frame f;
        f.type = h.type;
        f.seq = h.seq;
        f.src = h.src;
        f.dst = h.dst;
        f.data.assign(data, data + h.len);
        emit(TMK_FRAME_OK, consumed, true, h, data, h.len);
        frames.push_back(std::move(f));
        return long(consumed);
    }

    // capture ended: if the decoder is still waiting for samples, say so (side channel only)
    void finish() {
        if (decoding) {
            Header h{};
            emit(TMK_FRAME_TRUNCATED, 0, false, h, nullptr, 0);
        }
    }
};

}  // namespace orc
