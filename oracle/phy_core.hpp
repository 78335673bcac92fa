// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product; nothing under
// trackmaker-rs_b200/ may include, link or call this. Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs use it, as the checker / CPU baseline.
//
// CPU restatement (C++17) of the reference's src/phy line codes, CRC and frame format.
// Reference: HeZeBang/TrackMaker-rs @ e37d154. The reference is Rust and there is no
// rustc/cargo in this image, so oracle/_ref cannot be built; this restatement is pinned by
// the literals of the reference's own unit tests (see tests/test_oracle_pins.py) — the
// decoder state machine has NO reference test, so its parity is "pinned by restatement
// only" (SURVEY §8c).
//
// Build: g++ -O2 -mavx -ffp-contract=off -fno-fast-math (see oracle/Makefile).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

namespace orc {

// ---------------------------------------------------------------- crc.rs:4-75
// CRC-8, poly 0x07, init 0, MSB first, no reflection, no xor-out (crc.rs:7-22).
inline uint8_t crc8(const uint8_t* d, size_t n) {
    uint8_t c = 0;
    for (size_t i = 0; i < n; ++i) {
        c ^= d[i];
        for (int k = 0; k < 8; ++k) c = (c & 0x80) ? uint8_t((c << 1) ^ 0x07) : uint8_t(c << 1);
    }
    return c;
}

// crc.rs:54-60 — MSB-first expansion, one bit per byte.
inline std::vector<uint8_t> bytes_to_bits(const uint8_t* b, size_t n) {
    std::vector<uint8_t> bits;
    bits.reserve(n * 8);
    for (size_t i = 0; i < n; ++i)
        for (int k = 7; k >= 0; --k) bits.push_back((b[i] >> k) & 1);
    return bits;
}

// crc.rs:63-75 — trailing partial byte is zero padded in its low bits.
inline std::vector<uint8_t> bits_to_bytes(const uint8_t* bits, size_t n) {
    std::vector<uint8_t> out((n + 7) / 8, 0);
    for (size_t i = 0; i < n; ++i)
        if (bits[i]) out[i >> 3] |= uint8_t(0x80u >> (i & 7));
    return out;
}

// ---------------------------------------------------------------- line_coding.rs
struct LineCode {  // trait LineCode, line_coding.rs:6-24
    size_t spl;
    explicit LineCode(size_t s) : spl(s) {}
    virtual ~LineCode() = default;
    virtual std::vector<float> encode(const std::vector<uint8_t>& bits) const = 0;
    virtual std::vector<uint8_t> decode(const float* s, size_t n) const = 0;
    virtual size_t samples_for_bits(size_t nbits) const = 0;
    // default method, line_coding.rs:13-21: (bytes-1) x 0x33 then 0x5A, line coded
    std::vector<float> generate_preamble(size_t pattern_bytes) const {
        static const uint8_t k33[8] = {0, 0, 1, 1, 0, 0, 1, 1};
        static const uint8_t k5a[8] = {0, 1, 0, 1, 1, 0, 1, 0};
        std::vector<uint8_t> bits;
        for (size_t i = 0; i + 1 < pattern_bytes; ++i) bits.insert(bits.end(), k33, k33 + 8);
        bits.insert(bits.end(), k5a, k5a + 8);
        return encode(bits);
    }
};

// f32 mean of `spl` consecutive samples: sequential sum from 0 then one divide
// (line_coding.rs:104-113, 258-261).
inline float level_mean(const float* s, size_t spl) {
    float acc = 0.0f;
    for (size_t i = 0; i < spl; ++i) acc = acc + s[i];
    return acc / float(spl);
}

struct Manchester final : LineCode {  // line_coding.rs:63-132
    using LineCode::LineCode;
    std::vector<float> encode(const std::vector<uint8_t>& bits) const override {
        std::vector<float> out;
        out.reserve(bits.size() * spl * 2);
        for (uint8_t b : bits) {
            float first = b == 0 ? 1.0f : -1.0f;  // 0 -> high,low ; 1 -> low,high
            out.insert(out.end(), spl, first);
            out.insert(out.end(), spl, -first);
        }
        return out;
    }
    std::vector<uint8_t> decode(const float* s, size_t n) const override {
        size_t per_bit = spl * 2, nb = n / per_bit;
        std::vector<uint8_t> bits(nb);
        for (size_t i = 0; i < nb; ++i) {
            float a = level_mean(s + i * per_bit, spl);
            float b = level_mean(s + i * per_bit + spl, spl);
            bits[i] = a > b ? 0 : 1;  // tie decodes as 1
        }
        return bits;
    }
    size_t samples_for_bits(size_t nbits) const override { return nbits * spl * 2; }
};

// line_coding.rs:138-155
static const uint8_t k4b5b[16] = {0b11110, 0b01001, 0b10100, 0b10101, 0b01010, 0b01011, 0b01110, 0b01111,
                                  0b10010, 0b10011, 0b10110, 0b10111, 0b11010, 0b11011, 0b11100, 0b11101};

inline int decode_5b(uint8_t sym) {  // line_coding.rs:157-180; -1 = invalid symbol
    for (int i = 0; i < 16; ++i)
        if (k4b5b[i] == sym) return i;
    return -1;
}

struct FourBFiveB final : LineCode {  // line_coding.rs:182-333
    using LineCode::LineCode;
    std::vector<float> encode(const std::vector<uint8_t>& bits) const override {
        size_t nn = (bits.size() + 3) / 4;
        std::vector<float> out;
        out.reserve(nn * 5 * spl);
        float level = 1.0f;  // last_level is 1.0 on every call (&self), line_coding.rs:228
        for (size_t i = 0; i < nn; ++i) {
            unsigned nib = 0;
            for (size_t j = 0; j < 4 && i * 4 + j < bits.size(); ++j)
                if (bits[i * 4 + j]) nib |= 1u << (3 - j);  // last nibble zero padded
            uint8_t sym = k4b5b[nib];
            for (int j = 4; j >= 0; --j) {
                if ((sym >> j) & 1) level = -level;  // NRZI: 1 toggles
                out.insert(out.end(), spl, level);
            }
        }
        return out;
    }
    std::vector<uint8_t> decode(const float* s, size_t n) const override {
        std::vector<uint8_t> bits;
        if (n == 0) return bits;
        size_t nlev = n / spl;
        float last = 1.0f;  // prev_level_avg restarts at 1.0 on every call, :251
        unsigned sym = 0, cnt = 0;
        for (size_t i = 0; i < nlev; ++i) {
            float cur = level_mean(s + i * spl, spl);
            unsigned bit = (last * cur < 0.0f) ? 1u : 0u;
            if (std::fabs(cur) > 1e-6f) last = cur;  // keep the last non-zero level
            sym = (sym << 1) | bit;
            if (++cnt == 5) {
                int nib = decode_5b(uint8_t(sym));
                if (nib < 0) break;  // first invalid symbol stops the decode, :281-286
                for (int j = 3; j >= 0; --j) bits.push_back((nib >> j) & 1);
                sym = 0;
                cnt = 0;
            }
        }
        return bits;
    }
    size_t samples_for_bits(size_t nbits) const override { return ((nbits + 3) / 4) * 5 * spl; }
};

inline std::unique_ptr<LineCode> make_line_code(int kind, size_t spl) {  // LineCodingKind::create :40-49
    if (kind == 0) return std::make_unique<Manchester>(spl);
    return std::make_unique<FourBFiveB>(spl);
}

// ---------------------------------------------------------------- frame.rs
constexpr size_t kHeaderBytes = 7;  // consts.rs:48

struct Frame {  // frame.rs:34-41
    uint8_t type = 1, seq = 0, src = 0, dst = 0;
    std::vector<uint8_t> data;
    // frame.rs:74-102: [len_hi, len_lo, crc8(data), type, seq, src, dst, data...]
    std::vector<uint8_t> to_bytes() const {
        std::vector<uint8_t> b;
        size_t len = data.size();
        b.push_back(uint8_t(len >> 8));
        b.push_back(uint8_t(len & 0xFF));
        b.push_back(crc8(data.data(), data.size()));
        b.push_back(type);
        b.push_back(seq);
        b.push_back(src);
        b.push_back(dst);
        b.insert(b.end(), data.begin(), data.end());
        return b;
    }
};

struct Header {
    size_t len;
    uint8_t crc, type, seq, src, dst;
};

// frame.rs:109-143: needs 7 bytes and type in {1,2}. `raw` receives the fields even when the
// type check fails (side channel for verdict records).
inline bool parse_header(const std::vector<uint8_t>& bytes, Header* h, bool* have7) {
    *have7 = bytes.size() >= kHeaderBytes;
    if (!*have7) return false;
    h->len = (size_t(bytes[0]) << 8) | bytes[1];
    h->crc = bytes[2];
    h->type = bytes[3];
    h->seq = bytes[4];
    h->src = bytes[5];
    h->dst = bytes[6];
    return h->type == 1 || h->type == 2;
}

}  // namespace orc
