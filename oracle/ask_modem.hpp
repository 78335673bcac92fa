// ORACLE — TEST INFRASTRUCTURE ONLY (see phy_core.hpp header).
// CPU restatement of the project-1 chirp + BPSK modem of the reference,
// examples/ask.rs::test_sender_receiver (:526-858; live duplicates :104-262 TX, :264-524 RX).
// It has no stored vectors in the reference (only a run-time text comparison, :817-857), so its
// parity is pinned by this restatement plus the loopback of assets/think-different.txt.
//
// All arithmetic is f32 in the reference's operation order; sin is libm sinf (Rust's f32::sin
// lowers to the platform libm on Linux).
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace orc_ask {

constexpr int kPre = 440;          // preamble samples                      ask.rs:586-605
constexpr int kSpb = 44;           // samples per bit                       ask.rs:615
constexpr int kBits = 108;         // 100 frame bits + 8 zero "CRC" bits    ask.rs:612-613
constexpr int kFrame = kSpb * kBits;  // 4752
constexpr float kPi = 3.14159265358979323846f;  // std::f32::consts::PI

// carrier[k] = sin(2 pi 10000 t[k]), t[k] = k as f32 / 48000.0            ask.rs:574-583
inline std::vector<float> carrier(size_t n) {
    std::vector<float> c(n);
    for (size_t k = 0; k < n; ++k) {
        const float t = float(k) / 48000.0f;
        c[k] = sinf(2.0f * kPi * 10000.0f * t);
    }
    return c;
}

// chirp 2 kHz -> 10 kHz -> 2 kHz, cumulative trapezoid in f32              ask.rs:586-605
inline std::vector<float> preamble() {
    std::vector<float> fp;
    for (int i = 0; i < 220; ++i) fp.push_back(2000.0f + (8000.0f * float(i) / 219.0f));
    for (int i = 0; i < 220; ++i) fp.push_back(10000.0f - (8000.0f * float(i) / 219.0f));
    std::vector<float> p;
    float omega = 0.0f;
    p.push_back(sinf(2.0f * kPi * fp[0] * (0.0f / 48000.0f)));
    for (int i = 1; i < kPre; ++i) {
        const float dt = float(i) / 48000.0f - float(i - 1) / 48000.0f;
        omega += kPi * (fp[i] + fp[i - 1]) * dt;
        p.push_back(sinf(omega));
    }
    return p;
}

// TX: per frame gap1 zeros ++ preamble ++ 108 bits x 44 samples of +-carrier[k] (k frame-local) ++ gap2 zeros
// bits: n_frames x 100 (one bit per byte); the 8 trailing zero bits are added here.  ask.rs:608-637
inline std::vector<float> modulate(const uint8_t* bits, uint32_t n_frames, const uint32_t* gap1, const uint32_t* gap2) {
    const std::vector<float> pre = preamble(), car = carrier(48000);
    std::vector<float> out;
    for (uint32_t f = 0; f < n_frames; ++f) {
        out.insert(out.end(), gap1[f], 0.0f);
        out.insert(out.end(), pre.begin(), pre.end());
        for (int j = 0; j < kBits; ++j) {
            const float amp = (j < 100 && bits[f * 100 + j] == 1) ? 1.0f : -1.0f;
            for (int k = j * kSpb; k < (j + 1) * kSpb; ++k) out.push_back(car[k] * amp);
        }
        out.insert(out.end(), gap2[f], 0.0f);
    }
    return out;
}

struct Rec {
    uint64_t start_index;
    float sync_power;
    uint32_t accepted;
    uint8_t id;
    uint8_t bits[14];
};

// RX, one pass over the capture                                            ask.rs:668-796
inline std::vector<Rec> demodulate(const float* rx, size_t n) {
    const std::vector<float> pre = preamble(), car = carrier(kFrame);
    std::vector<Rec> recs;
    float power = 0.0f;
    size_t start_index = 0;
    std::vector<float> fifo(kPre, 0.0f);
    size_t head = 0;  // fifo[(head + k) % 440] is sync_fifo[k] (rotate_left(1) + store at [439])
    float local_max = 0.0f;
    std::vector<float> dec;
    int state = 0;
    for (size_t i = 0; i < n; ++i) {
        const float x = rx[i];
        power = power * (1.0f - 1.0f / 64.0f) + x * x / 64.0f;  // :696-697
        if (state == 0) {
            fifo[head] = x;  // after the rotation the new sample sits at [439]
            head = (head + 1) % kPre;
            float sum = 0.0f;  // iter().zip().map(a*b).sum::<f32>()  :704-709
            for (int k = 0; k < kPre; ++k) sum = sum + fifo[(head + k) % kPre] * pre[k];
            const float sp = sum / 200.0f;
            if (sp > power * 2.0f && sp > local_max && sp > 0.05f) {  // :711-716
                local_max = sp;
                start_index = i;
            } else if (i > start_index + 200 && start_index != 0) {  // :717-723
                Rec r{};
                r.start_index = start_index;
                r.sync_power = local_max;
                recs.push_back(r);  // completed below when the frame has been gathered
                local_max = 0.0f;
                std::fill(fifo.begin(), fifo.end(), 0.0f);
                state = 1;
                dec.assign(rx + start_index + 1, rx + i);  // the sample at i itself is never appended
            }
        } else {
            dec.push_back(x);
            if (dec.size() == size_t(kFrame)) {
                std::vector<float> sm(kFrame);
                for (int j = 0; j < kFrame; ++j) {  // :733-747
                    const int a = j >= 5 ? j - 5 : 0, b = j + 6 < kFrame ? j + 6 : kFrame;
                    float s = 0.0f;
                    for (int k = a; k < b; ++k) s = s + dec[k] * car[k];
                    sm[j] = s / float(b - a);
                }
                Rec& r = recs.back();
                uint8_t id = 0;
                for (int j = 0; j < kBits; ++j) {  // :749-761
                    float s = 0.0f;
                    for (int k = 10 + j * kSpb; k < 30 + j * kSpb; ++k) s = s + sm[k];
                    const bool bit = s > 0.0f;
                    if (bit) r.bits[j >> 3] |= uint8_t(0x80u >> (j & 7));
                    if (j < 8 && bit) id = uint8_t(id + (1u << (7 - j)));
                }
                r.id = id;
                r.accepted = (id > 0 && id <= 100) ? 1u : 0u;  // :771
                start_index = 0;
                dec.clear();
                state = 0;
            }
        }
    }
    if (state == 1) recs.pop_back();  // capture ended while the frame was being gathered: nothing decoded
    return recs;
}

}  // namespace orc_ask
