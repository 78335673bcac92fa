// ORACLE — TEST INFRASTRUCTURE ONLY (see phy_core.hpp header).
// Self-oracle of the multi-carrier modem (FSK and OFDM demodulation on a batched FFT: north_star stages 3-4,
// BASELINE config 3). **PARITY UNPINNED BY THE REFERENCE**: the reference has no FSK / OFDM / FFT code at all
// (SURVEY §0, §8c last row); docs/proj/report.md:177-319 (amodem-style multi-carrier PSK: pilot prefix, training
// symbols for the equaliser, per-carrier nearest-neighbour decision, length prefix + checksum framing) is prose and
// serves as the structural guide only. The specification below is the build's own; this file is
// its plain, sequential CPU statement, the CUDA path (csrc/mcm.cu) is compared with it bit for bit.
//
// Frame on the air:  [440-sample chirp of examples/ask.rs:586-605][4752-sample body]   (the envelope of an ask.rs
// frame, so the reference's own chirp synchroniser — ask.rs:693-723, restated in ask_modem.hpp — locks on it:
// start_index = sample index of the sync peak, body sample b = rx[start_index + 1 + b]).
//   OFDM body: 16 zeros | training symbol | 13 data symbols | 256 zeros; a symbol is 64 samples of cyclic prefix +
//              256 samples; 32 carriers on FFT bins 16..47 (3.0 - 8.8 kHz); QPSK (+-1 +-j)/sqrt(2), 2 bits per
//              carrier and symbol; one-tap equaliser from the training symbol (known points 1, j, -1, -j).
//   FSK body:  16 zeros | 56 symbols | 256 zeros; a symbol is 16 + 64 samples of ONE of 16 tones (FFT bins 4..19,
//              3.0 - 14.25 kHz), 4 bits per symbol; decision = the strongest of the 16 bins (first wins a tie).
//   Bits -> bytes MSB first (crc.rs:41-60); bytes = [seq][len][payload ...][crc8(payload), crc.rs:7-22][zero padding];
//   verdicts: len beyond the capacity (OFDM 100, FSK 25) -> BAD_LEN, checksum mismatch -> CRC_FAIL, else OK.
// Arithmetic: f32, every operation rounded separately (no FMA), in exactly the order written here; tables are
// computed in double and rounded once. FFT: Stockham autosort radix-4, decimation in frequency, twiddle skipped for p = 0.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "ask_modem.hpp"
#include "phy_core.hpp"

namespace orc_mcm {

constexpr int kModeFsk = 0, kModeOfdm = 1;
constexpr int kBody = 4752, kLead = 16;
constexpr int kN = 256, kCp = 64, kCarriers = 32, kBin0 = 16, kDataSyms = 13;   // OFDM
constexpr int kNf = 64, kCpf = 16, kTones = 16, kToneBin0 = 4, kFskSyms = 56;  // FSK
constexpr int kMaxPayloadOfdm = 100, kMaxPayloadFsk = 25;
constexpr float kGainData = 0.088388346f;  // 0.125 / sqrt(2): the +-1 +-j components of a data point
constexpr float kGainTrain = 0.125f;
constexpr float kGainFsk = 0.70710678f;

struct Cpx { float re, im; };

// W[t] = exp(-2 pi i t / n), t = 0 .. n-1 (double, rounded once)
inline std::vector<Cpx> twiddles(int n) {
    std::vector<Cpx> w(n);
    for (int t = 0; t < n; ++t) {
        const double a = 6.283185307179586476925286766559 * double(t) / double(n);
        w[t].re = float(std::cos(a));
        w[t].im = float(-std::sin(a));
    }
    return w;
}
// training points: index 0..3 -> 1, j, -1, -j from a 32-bit LCG
inline void training(uint8_t t[kCarriers]) {
    uint32_t st = 0x1234u;
    for (int i = 0; i < kCarriers; ++i) {
        st = st * 1103515245u + 12345u;
        t[i] = uint8_t((st >> 16) & 3u);
    }
}

inline Cpx cmul(Cpx w, Cpx v) {
    Cpx r;
    r.re = w.re * v.re - w.im * v.im;  // (-ffp-contract=off: three roundings)
    r.im = w.re * v.im + w.im * v.re;
    return r;
}

// in-order complex FFT of n = 4^k points (n = 64 or 256). x is overwritten with scratch; the result is returned.
inline std::vector<Cpx> fft(std::vector<Cpx> x, const std::vector<Cpx>& W) {
    const int N = int(x.size());
    std::vector<Cpx> y(N);
    Cpx* src = x.data();
    Cpx* dst = y.data();
    for (int n = N, s = 1; n >= 4; n /= 4, s *= 4) {
        const int n1 = n / 4;
        for (int p = 0; p < n1; ++p) {
            const Cpx w1 = W[size_t(p) * (N / n)], w2 = W[size_t(2 * p) * (N / n)], w3 = W[size_t(3 * p) * (N / n)];
            for (int q = 0; q < s; ++q) {
                const Cpx a = src[q + s * p], b = src[q + s * (p + n1)], c = src[q + s * (p + 2 * n1)], d = src[q + s * (p + 3 * n1)];
                const Cpx apc{a.re + c.re, a.im + c.im}, amc{a.re - c.re, a.im - c.im};
                const Cpx bpd{b.re + d.re, b.im + d.im}, bmd{b.re - d.re, b.im - d.im};
                const Cpx y0{apc.re + bpd.re, apc.im + bpd.im};
                const Cpx y1{amc.re + bmd.im, amc.im - bmd.re};  // amc - j (b - d)
                const Cpx y2{apc.re - bpd.re, apc.im - bpd.im};
                const Cpx y3{amc.re - bmd.im, amc.im + bmd.re};  // amc + j (b - d)
                dst[q + s * (4 * p)] = y0;
                dst[q + s * (4 * p + 1)] = p ? cmul(w1, y1) : y1;
                dst[q + s * (4 * p + 2)] = p ? cmul(w2, y2) : y2;
                dst[q + s * (4 * p + 3)] = p ? cmul(w3, y3) : y3;
            }
        }
        Cpx* t = src;
        src = dst;
        dst = t;
    }
    return std::vector<Cpx>(src, src + N);
}

inline std::vector<Cpx> fft_real(const float* v, int n, const std::vector<Cpx>& W) {
    std::vector<Cpx> x(n);
    for (int i = 0; i < n; ++i) x[i] = Cpx{v[i], 0.0f};
    return fft(x, W);
}

inline int capacity(int mode) { return mode == kModeOfdm ? kMaxPayloadOfdm : kMaxPayloadFsk; }
inline int frame_bytes(int mode) { return mode == kModeOfdm ? kDataSyms * kCarriers * 2 / 8 : kFskSyms * 4 / 8; }  // 104 / 28

// the bytes a frame carries
inline std::vector<uint8_t> frame_to_bytes(int mode, uint8_t seq, const uint8_t* payload, uint32_t len) {
    std::vector<uint8_t> b(frame_bytes(mode), 0);
    b[0] = seq;
    b[1] = uint8_t(len);
    for (uint32_t i = 0; i < len; ++i) b[2 + i] = payload[i];
    b[2 + len] = orc::crc8(payload, len);
    return b;
}

// 4752 body samples of one frame
inline std::vector<float> modulate_body(int mode, uint8_t seq, const uint8_t* payload, uint32_t len) {
    const std::vector<uint8_t> bytes = frame_to_bytes(mode, seq, payload, len);
    auto bit = [&](int j) { return (bytes[j >> 3] >> (7 - (j & 7))) & 1; };
    std::vector<float> out(kBody, 0.0f);
    if (mode == kModeOfdm) {
        std::vector<float> C(kN), S(kN);
        for (int m = 0; m < kN; ++m) {
            const double a = 6.283185307179586476925286766559 * double(m) / double(kN);
            C[m] = float(std::cos(a));
            S[m] = float(std::sin(a));
        }
        uint8_t tr[kCarriers];
        training(tr);
        for (int sym = 0; sym <= kDataSyms; ++sym) {
            float sr[kCarriers], si[kCarriers];
            for (int i = 0; i < kCarriers; ++i) {
                if (sym == 0) {
                    const float re[4] = {1.f, 0.f, -1.f, 0.f}, im[4] = {0.f, 1.f, 0.f, -1.f};
                    sr[i] = kGainTrain * re[tr[i]];
                    si[i] = kGainTrain * im[tr[i]];
                } else {
                    const int j = ((sym - 1) * kCarriers + i) * 2;
                    sr[i] = bit(j) ? -kGainData : kGainData;
                    si[i] = bit(j + 1) ? -kGainData : kGainData;
                }
            }
            float* o = out.data() + kLead + sym * (kN + kCp);
            for (int n = 0; n < kN + kCp; ++n) {
                const int t = (n + kN - kCp) % kN;  // cyclic prefix: the last 64 samples first
                float acc = 0.0f;
                for (int i = 0; i < kCarriers; ++i) {
                    const int m = ((kBin0 + i) * t) % kN;
                    acc = acc + sr[i] * C[m];
                    acc = acc - si[i] * S[m];
                }
                o[n] = acc;
            }
        }
    } else {
        std::vector<float> S(kNf);
        for (int m = 0; m < kNf; ++m) S[m] = float(std::sin(6.283185307179586476925286766559 * double(m) / double(kNf)));
        for (int sym = 0; sym < kFskSyms; ++sym) {
            const int tone = (bit(4 * sym) << 3) | (bit(4 * sym + 1) << 2) | (bit(4 * sym + 2) << 1) | bit(4 * sym + 3);
            float* o = out.data() + kLead + sym * (kNf + kCpf);
            for (int n = 0; n < kNf + kCpf; ++n) {
                const int t = (n + kNf - kCpf) % kNf;
                o[n] = kGainFsk * S[((kToneBin0 + tone) * t) % kNf];
            }
        }
    }
    return out;
}

// TX: per frame gap1 zeros ++ chirp ++ body ++ gap2 zeros
inline std::vector<float> modulate(int mode, const uint8_t* payload, const uint32_t* off, const uint8_t* seq, uint32_t n_frames,
                                   const uint32_t* gap1, const uint32_t* gap2) {
    const std::vector<float> pre = orc_ask::preamble();
    std::vector<float> out;
    for (uint32_t f = 0; f < n_frames; ++f) {
        out.insert(out.end(), gap1[f], 0.0f);
        out.insert(out.end(), pre.begin(), pre.end());
        const std::vector<float> body = modulate_body(mode, seq[f], payload + off[f], off[f + 1] - off[f]);
        out.insert(out.end(), body.begin(), body.end());
        out.insert(out.end(), gap2[f], 0.0f);
    }
    return out;
}

struct Rec {
    uint64_t start_index;
    float sync_power;
    uint32_t status;  // TMK_FRAME_OK / TMK_FRAME_BAD_LEN / TMK_FRAME_CRC_FAIL
    uint8_t seq, len, crc;
    std::vector<uint8_t> payload;
};

// the 104 / 28 bytes decided from the body that starts at rx[s + 1]
inline std::vector<uint8_t> demod_body(int mode, const float* rx, uint64_t s) {
    std::vector<uint8_t> bytes(frame_bytes(mode), 0);
    auto put = [&](int j, int b) { if (b) bytes[j >> 3] |= uint8_t(0x80u >> (j & 7)); };
    const float* body = rx + s + 1;
    if (mode == kModeOfdm) {
        const std::vector<Cpx> W = twiddles(kN);
        uint8_t tr[kCarriers];
        training(tr);
        const std::vector<Cpx> R = fft_real(body + kLead + kCp, kN, W);
        for (int sym = 1; sym <= kDataSyms; ++sym) {
            const std::vector<Cpx> Y = fft_real(body + kLead + sym * (kN + kCp) + kCp, kN, W);
            for (int i = 0; i < kCarriers; ++i) {
                const Cpx y = Y[kBin0 + i], r = R[kBin0 + i];
                const float zr = y.re * r.re + y.im * r.im;  // y conj(r)
                const float zi = y.im * r.re - y.re * r.im;
                float er, ei;                              // times the training point
                switch (tr[i]) {
                    case 0: er = zr; ei = zi; break;
                    case 1: er = -zi; ei = zr; break;
                    case 2: er = -zr; ei = -zi; break;
                    default: er = zi; ei = -zr; break;
                }
                const int j = ((sym - 1) * kCarriers + i) * 2;
                put(j, er < 0.0f);
                put(j + 1, ei < 0.0f);
            }
        }
    } else {
        const std::vector<Cpx> W = twiddles(kNf);
        for (int sym = 0; sym < kFskSyms; ++sym) {
            const std::vector<Cpx> X = fft_real(body + kLead + sym * (kNf + kCpf) + kCpf, kNf, W);
            int best = 0;
            float eb = 0.0f;
            for (int m = 0; m < kTones; ++m) {
                const Cpx v = X[kToneBin0 + m];
                const float e = v.re * v.re + v.im * v.im;
                if (m == 0 || e > eb) {  // the first strongest tone (a NaN never wins, a NaN in tone 0 never loses)
                    eb = e;
                    best = m;
                }
            }
            for (int k = 0; k < 4; ++k) put(4 * sym + k, (best >> (3 - k)) & 1);
        }
    }
    return bytes;
}

// RX of one capture: the chirp synchroniser of ask.rs finds the frames, the body of each is demodulated
inline std::vector<Rec> demodulate(int mode, const float* rx, size_t n) {
    std::vector<Rec> out;
    for (const orc_ask::Rec& a : orc_ask::demodulate(rx, n)) {
        Rec r;
        r.start_index = a.start_index;
        r.sync_power = a.sync_power;
        const std::vector<uint8_t> bytes = demod_body(mode, rx, a.start_index);
        r.seq = bytes[0];
        r.len = bytes[1];
        r.crc = 0;
        if (int(r.len) > capacity(mode)) {
            r.status = 2;  // TMK_FRAME_BAD_LEN
        } else {
            r.crc = bytes[2 + r.len];
            r.payload.assign(bytes.begin() + 2, bytes.begin() + 2 + r.len);
            r.status = orc::crc8(r.payload.data(), r.payload.size()) == r.crc ? 0u : 5u;  // OK / CRC_FAIL
        }
        out.push_back(r);
    }
    return out;
}

}  // namespace orc_mcm
