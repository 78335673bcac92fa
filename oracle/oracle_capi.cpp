// ORACLE — TEST INFRASTRUCTURE ONLY (see phy_core.hpp header). C entry points for ctypes.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>

#include "ask_modem.hpp"
#include "mcm_modem.hpp"
#include "ip_fragments.hpp"
#include "phy_decoder.hpp"

using namespace orc;

extern "C" {

uint8_t orc_crc8(const uint8_t* d, uint64_t n) { return crc8(d, n); }

uint64_t orc_bytes_to_bits(const uint8_t* b, uint64_t n, uint8_t* bits) {
    auto v = bytes_to_bits(b, n);
    std::copy(v.begin(), v.end(), bits);
    return v.size();
}

uint64_t orc_bits_to_bytes(const uint8_t* bits, uint64_t n, uint8_t* bytes) {
    auto v = bits_to_bytes(bits, n);
    std::copy(v.begin(), v.end(), bytes);
    return v.size();
}

uint64_t orc_samples_for_bits(int kind, uint32_t spl, uint64_t nbits) {
    return make_line_code(kind, spl)->samples_for_bits(nbits);
}

// LineCode::encode ; returns number of samples (writes at most cap)
uint64_t orc_line_encode(int kind, uint32_t spl, const uint8_t* bits, uint64_t nbits, float* out, uint64_t cap) {
    auto v = make_line_code(kind, spl)->encode(std::vector<uint8_t>(bits, bits + nbits));
    std::copy(v.begin(), v.begin() + std::min<uint64_t>(cap, v.size()), out);
    return v.size();
}

uint64_t orc_line_decode(int kind, uint32_t spl, const float* s, uint64_t n, uint8_t* bits, uint64_t cap) {
    auto v = make_line_code(kind, spl)->decode(s, n);
    std::copy(v.begin(), v.begin() + std::min<uint64_t>(cap, v.size()), bits);
    return v.size();
}

uint64_t orc_preamble(int kind, uint32_t spl, uint32_t pbytes, float* out, uint64_t cap) {
    auto v = make_line_code(kind, spl)->generate_preamble(pbytes);
    std::copy(v.begin(), v.begin() + std::min<uint64_t>(cap, v.size()), out);
    return v.size();
}

// Frame::to_bytes
uint64_t orc_frame_to_bytes(const tmk_frame_desc* f, const uint8_t* payload, uint8_t* out) {
    Frame fr;
    fr.type = f->type;
    fr.seq = f->seq;
    fr.src = f->src;
    fr.dst = f->dst;
    fr.data.assign(payload + f->payload_off, payload + f->payload_off + f->len);
    auto v = fr.to_bytes();
    std::copy(v.begin(), v.end(), out);
    return v.size();
}

// PhyEncoder::encode_frames
uint64_t orc_modulate(const tmk_phy_config* cfg, const tmk_frame_desc* frames, uint32_t n_frames,
                      const uint8_t* payload, uint32_t gap, float* out, uint64_t cap) {
    Encoder enc(cfg->spl, cfg->preamble_bytes, cfg->line_code);
    std::vector<Frame> fs(n_frames);
    for (uint32_t i = 0; i < n_frames; ++i) {
        fs[i].type = frames[i].type;
        fs[i].seq = frames[i].seq;
        fs[i].src = frames[i].src;
        fs[i].dst = frames[i].dst;
        fs[i].data.assign(payload + frames[i].payload_off, payload + frames[i].payload_off + frames[i].len);
    }
    auto v = enc.encode_frames(fs, gap);
    std::copy(v.begin(), v.begin() + std::min<uint64_t>(cap, v.size()), out);
    return v.size();
}

static void decode_one(const tmk_phy_config& cfg, const float* s, uint64_t n, uint32_t stream_id,
                       std::vector<tmk_frame_rec>& recs, std::vector<uint8_t>& payload) {
    Decoder d(cfg);
    d.stream_id = stream_id;
    d.recs = &recs;
    d.payload = &payload;
    if (cfg.chunk == 0 || cfg.chunk >= n) {
        d.process_samples(s, n);
    } else {
        for (uint64_t p = 0; p < n; p += cfg.chunk) d.process_samples(s + p, std::min<uint64_t>(cfg.chunk, n - p));
    }
    d.finish();
}

// One capture through a fresh PhyDecoder, fed in cfg->chunk sized process_samples calls.
// Returns the number of records (writes at most rec_cap); *payload_len receives payload bytes.
uint64_t orc_decode_stream(const tmk_phy_config* cfg, const float* s, uint64_t n, uint32_t stream_id,
                           tmk_frame_rec* recs, uint64_t rec_cap, uint8_t* payload, uint64_t payload_cap,
                           uint64_t* payload_len) {
    std::vector<tmk_frame_rec> r;
    std::vector<uint8_t> p;
    decode_one(*cfg, s, n, stream_id, r, p);
    for (size_t i = 0; i < r.size() && i < rec_cap; ++i) recs[i] = r[i];
    std::copy(p.begin(), p.begin() + std::min<uint64_t>(payload_cap, p.size()), payload);
    if (payload_len) *payload_len = p.size();
    return r.size();
}

// One capture through ONE PhyDecoder fed with process_samples calls of the given sizes (the live cadence,
// src/mac/csma.rs:446-465). feed_of_rec[i] receives the index of the call that returned record i. The
// capture is not finished: a frame still being gathered when the feeds end yields no record.
uint64_t orc_decode_feeds(const tmk_phy_config* cfg, const float* s, uint64_t n, const uint64_t* feed_sizes,
                          uint32_t n_feeds, uint32_t stream_id, tmk_frame_rec* recs, uint32_t* feed_of_rec,
                          uint64_t rec_cap, uint8_t* payload, uint64_t payload_cap, uint64_t* payload_len) {
    std::vector<tmk_frame_rec> r;
    std::vector<uint8_t> p;
    Decoder d(*cfg);
    d.stream_id = stream_id;
    d.recs = &r;
    d.payload = &p;
    uint64_t pos = 0, done = 0;
    for (uint32_t f = 0; f < n_feeds; ++f) {
        const uint64_t len = std::min<uint64_t>(feed_sizes[f], n - pos);
        d.process_samples(s + pos, len);
        pos += len;
        for (; done < r.size(); ++done)
            if (feed_of_rec && done < rec_cap) feed_of_rec[done] = f;
    }
    for (size_t i = 0; i < r.size() && i < rec_cap; ++i) recs[i] = r[i];
    std::copy(p.begin(), p.begin() + std::min<uint64_t>(payload_cap, p.size()), payload);
    if (payload_len) *payload_len = p.size();
    return r.size();
}

// Ragged batch over a thread pool (one PhyDecoder per stream, like one decoder per MAC node,
// src/mac/csma.rs:37-47). Records come back sorted by stream; payload offsets are global.
// Returns records; *seconds receives the wall time of the decode region.
uint64_t orc_decode_batch_mt(const float* samples, const tmk_stream* streams, uint32_t n_streams, int n_threads,
                             tmk_frame_rec* recs, uint64_t rec_cap, uint8_t* payload, uint64_t payload_cap,
                             uint64_t* payload_len, double* seconds) {
    if (n_threads < 1) n_threads = 1;
    std::vector<std::vector<tmk_frame_rec>> R(n_streams);
    std::vector<std::vector<uint8_t>> Pl(n_streams);
    std::atomic<uint32_t> next{0};
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&] {
        for (;;) {
            uint32_t s = next.fetch_add(1);
            if (s >= n_streams) return;
            decode_one(streams[s].cfg, samples + streams[s].offset, streams[s].n_samples, s, R[s], Pl[s]);
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < n_threads; ++t) th.emplace_back(work);
    work();
    for (auto& t : th) t.join();
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    uint64_t nr = 0, np = 0;
    for (uint32_t s = 0; s < n_streams; ++s) {
        for (auto r : R[s]) {
            r.payload_off += np;
            if (recs && nr < rec_cap) recs[nr] = r;
            ++nr;
        }
        if (payload && np + Pl[s].size() <= payload_cap) std::copy(Pl[s].begin(), Pl[s].end(), payload + np);
        np += Pl[s].size();
    }
    if (payload_len) *payload_len = np;
    return nr;
}

// The rolling window energy of search_for_preamble (decoder.rs:133-136, 224-233) entered at `seed`, at the given
// ascending windows: fresh sequential seed sum, then  e = e - x[i]^2 + x[i+P]^2  with the clamp at 0.
void orc_rolling_energy(const float* s, uint64_t n, uint32_t P, uint64_t seed, const uint64_t* windows, uint32_t n_windows,
                        float* out) {
    float e = seq_energy(s + seed, P);
    uint64_t pos = seed;
    for (uint32_t k = 0; k < n_windows; ++k) {
        for (; pos < windows[k]; ++pos) {
            float leaving = s[pos], entering = s[pos + P];
            e = e - leaving * leaving + entering * entering;
            if (e < 0.0f) e = 0.0f;
        }
        out[k] = e;
    }
    (void)n;
}

// Correlation of every window in reference arithmetic (one search_for_preamble pass that
// never stops: seed at window 0, rolling energy, AVX-order dot, energy gate).
uint64_t orc_correlate(const tmk_phy_config* cfg, const float* s, uint64_t n, float* out) {
    Decoder d(*cfg);
    size_t P = d.preamble.size();
    if (n < P) return 0;
    size_t windows = n - P + 1;
    float energy = seq_energy(s, P);
    for (size_t i = 0; i < windows; ++i) {
        out[i] = energy < 1e-6f ? 0.0f : dot_avx_order(s + i, d.preamble.data(), P) / (std::sqrt(energy) * d.preamble_energy);
        if (i + 1 < windows) {
            energy = energy - s[i] * s[i] + s[i + P] * s[i + P];
            if (energy < 0.0f) energy = 0.0f;
        }
    }
    return windows;
}

// mac::is_channel_busy (src/mac/mod.rs:18-27): -1 = None
int orc_channel_busy(const float* s, uint64_t n) {
    if (n < 20) return -1;
    for (uint64_t i = 0; i < n; ++i)
        if (std::fabs(s[i]) > 0.5f) return 1;
    return 0;
}

// ---- chirp + BPSK modem (examples/ask.rs)
void orc_ask_tables(float* preamble440, float* carrier, uint64_t n_carrier) {
    auto p = orc_ask::preamble();
    std::copy(p.begin(), p.end(), preamble440);
    auto c = orc_ask::carrier(n_carrier);
    std::copy(c.begin(), c.end(), carrier);
}

uint64_t orc_ask_modulate(const uint8_t* bits, uint32_t n_frames, const uint32_t* gap1, const uint32_t* gap2,
                          float* out, uint64_t cap) {
    auto v = orc_ask::modulate(bits, n_frames, gap1, gap2);
    std::copy(v.begin(), v.begin() + std::min<uint64_t>(cap, v.size()), out);
    return v.size();
}

// recs: tmk_ask_rec layout (include/tmkphy.h)
uint64_t orc_ask_demodulate(const float* rx, uint64_t n, uint32_t stream, tmk_ask_rec* recs, uint64_t cap) {
    auto v = orc_ask::demodulate(rx, n);
    for (size_t i = 0; i < v.size() && i < cap; ++i) {
        tmk_ask_rec r{};
        r.stream = stream;
        r.accepted = v[i].accepted;
        r.start_index = v[i].start_index;
        r.sync_power = v[i].sync_power;
        r.id = v[i].id;
        std::memcpy(r.bits, v[i].bits, 14);
        recs[i] = r;
    }
    return v.size();
}

// ---- multi-carrier modem (build-defined FSK / OFDM spec, mcm_modem.hpp; parity unpinned by the reference)
uint64_t orc_mcm_modulate(int mode, const uint8_t* payload, const uint32_t* off, const uint8_t* seq, uint32_t n_frames,
                          const uint32_t* gap1, const uint32_t* gap2, float* out, uint64_t cap) {
    auto v = orc_mcm::modulate(mode, payload, off, seq, n_frames, gap1, gap2);
    std::copy(v.begin(), v.begin() + std::min<uint64_t>(cap, v.size()), out);
    return v.size();
}
// recs: tmk_mcm_rec layout; payload bytes appended at *payload_len
uint64_t orc_mcm_demodulate(int mode, const float* rx, uint64_t n, uint32_t stream, tmk_mcm_rec* recs, uint64_t cap,
                            uint8_t* payload, uint64_t payload_cap, uint64_t* payload_len) {
    auto v = orc_mcm::demodulate(mode, rx, n);
    uint64_t pl = *payload_len;
    for (size_t i = 0; i < v.size(); ++i) {
        if (i < cap) {
            tmk_mcm_rec r{};
            r.stream = stream;
            r.status = v[i].status;
            r.start_index = v[i].start_index;
            r.payload_off = pl;
            r.sync_power = v[i].sync_power;
            r.seq = v[i].seq;
            r.len = v[i].len;
            r.crc = v[i].crc;
            r.mode = uint8_t(mode);
            recs[i] = r;
        }
        for (uint8_t b : v[i].payload) {
            if (pl < payload_cap) payload[pl] = b;
            ++pl;
        }
    }
    *payload_len = pl;
    return v.size();
}
// complex FFT of n_ffts real rows of n (64 | 256) samples: out = n_ffts x n x (re, im)
void orc_mcm_fft(int n, const float* in, uint32_t n_ffts, float* out) {
    const auto W = orc_mcm::twiddles(n);
    for (uint32_t f = 0; f < n_ffts; ++f) {
        const auto X = orc_mcm::fft_real(in + size_t(f) * n, n, W);
        for (int k = 0; k < n; ++k) {
            out[(size_t(f) * n + k) * 2] = X[k].re;
            out[(size_t(f) * n + k) * 2 + 1] = X[k].im;
        }
    }
}

// ---- 16-bit PCM (dump.rs:37-40, codec.rs:121-125)
void orc_pcm16_encode(const float* s, uint64_t n, int16_t* out) {
    for (uint64_t i = 0; i < n; ++i) out[i] = pcm16_of(s[i]);
}
void orc_pcm16_to_f32(const int16_t* p, uint64_t n, float* out) {
    for (uint64_t i = 0; i < n; ++i) out[i] = f32_of_pcm16(p[i]);
}

// ---- IP fragments (fragmentation.rs)
uint16_t orc_fraginfo_to_u16(uint16_t id, int more, uint16_t off) { return FragmentationInfo{id, more != 0, off}.to_u16(); }
void orc_fraginfo_from_u16(uint16_t v, int* more, uint16_t* off) {
    const auto fi = FragmentationInfo::from_u16(v);
    *more = fi.more_fragments;
    *off = fi.fragment_offset;
}

// one IpFragmenter over a list of packets (packet i = data[off[i] .. off[i+1])). Records as tmk_fragment_batch.
uint64_t orc_fragment_batch(const uint8_t* data, const uint64_t* off, uint32_t n, uint32_t mtu, uint16_t first_id,
                            tmk_packet_rec* frags, uint64_t frag_cap, uint8_t* bytes, uint64_t byte_cap, uint64_t* n_bytes,
                            uint16_t* next_id) {
    IpFragmenter fr(mtu);
    fr.next_id = first_id;
    uint64_t nf = 0, nb = 0;
    for (uint32_t i = 0; i < n; ++i) {
        Bytes pkt(data + off[i], data + off[i + 1]);
        std::vector<Bytes> out;
        const uint16_t id = fr.next_id;
        const std::string err = fr.fragment_packet(pkt, out);
        if (!err.empty()) {
            const uint32_t st = err.find("too small for header") != std::string::npos ? TMK_PKT_E_SHORT
                                : err.find("header length") != std::string::npos    ? TMK_PKT_E_IHL
                                                                                     : TMK_PKT_E_MTU;
            if (nf < frag_cap) frags[nf] = tmk_packet_rec{i, st, 0, 0, nb, 0, 0, 0};
            ++nf;
            continue;
        }
        const bool cut = fr.next_id != id;
        for (size_t f = 0; f < out.size(); ++f) {
            if (nf < frag_cap) frags[nf] = tmk_packet_rec{i, TMK_PKT_OK, uint32_t(f), uint32_t(out[f].size()), nb,
                                                          uint16_t(cut ? id : 0), uint16_t(out.size()), 0};
            ++nf;
            if (nb + out[f].size() <= byte_cap) std::copy(out[f].begin(), out[f].end(), bytes + nb);
            nb += out[f].size();
        }
    }
    *n_bytes = nb;
    *next_id = fr.next_id;
    return nf;
}

// per link one IpReassembler fed with the frames receive_packet would hand it (acoustic_interface.rs:343-366);
// records sorted by stream as tmk_decode_batch returns them.
uint64_t orc_reassemble_batch(const tmk_frame_rec* recs, uint64_t n_recs, const uint8_t* payload, uint32_t n_links,
                              tmk_packet_rec* pkts, uint64_t pkt_cap, uint8_t* bytes, uint64_t byte_cap, uint64_t* n_bytes) {
    uint64_t np = 0, nb = 0, i = 0;
    while (i < n_recs) {
        const uint32_t link = recs[i].stream;
        IpReassembler ra;
        for (; i < n_recs && recs[i].stream == link; ++i) {
            const tmk_frame_rec& r = recs[i];
            if (link >= n_links || r.status != TMK_FRAME_OK) continue;
            if (!(r.type == TMK_FRAME_TYPE_DATA || (r.type == TMK_FRAME_TYPE_ACK && r.len != 0))) continue;
            Bytes out;
            const uint8_t* p = payload + r.payload_off;
            const int rc = ra.process_fragment(p, r.len, out);
            if (rc == 0) continue;
            tmk_packet_rec pr{link, TMK_PKT_OK, uint32_t(i), 0, nb, 0, 0, 0};
            if (rc < 0) {
                pr.status = rc == -1 ? TMK_PKT_E_SHORT : TMK_PKT_E_IHL;
            } else {
                pr.len = uint32_t(out.size());
                pr.identification = uint16_t(p[4] << 8 | p[5]);
                pr.n_fragments = uint16_t(ra.joined);
                if (nb + out.size() <= byte_cap) std::copy(out.begin(), out.end(), bytes + nb);
                nb += out.size();
            }
            if (np < pkt_cap) pkts[np] = pr;
            ++np;
        }
    }
    *n_bytes = nb;
    return np;
}

int orc_hardware_threads(void) { return int(std::thread::hardware_concurrency()); }

}  // extern "C"
