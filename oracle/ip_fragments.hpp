// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference's payload layer above the PHY and of its
// 16-bit PCM conversions; never linked into libtmkphy.
//   FragmentationInfo      src/net/fragmentation.rs:6-57
//   IpFragmenter           src/net/fragmentation.rs:60-204
//   IpReassembler          src/net/fragmentation.rs:207-401
//   frame filter           src/mac/acoustic_interface.rs:343-366 (receive_packet)
//   pcm16                  src/utils/dump.rs:37-40, src/audio/codec.rs:121-125
// Pinned by the reference's own unit tests (fragmentation.rs:404-515), restated in tests/test_oracle_pins.py.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <map>
#include <string>
#include <utility>
#include <vector>

namespace orc {

struct FragmentationInfo {  // fragmentation.rs:6-57
    uint16_t identification;
    bool more_fragments;
    uint16_t fragment_offset;
    uint16_t to_u16() const {  // :32-45
        uint16_t value = 0;
        if (more_fragments) value |= 0x2000;
        value |= fragment_offset & 0x1FFF;
        return value;
    }
    static FragmentationInfo from_u16(uint16_t value) {  // :47-57
        return FragmentationInfo{0, (value & 0x2000) != 0, uint16_t(value & 0x1FFF)};
    }
};

using Bytes = std::vector<uint8_t>;

struct IpFragmenter {  // fragmentation.rs:60-204
    size_t mtu;
    uint16_t next_id = 0;
    explicit IpFragmenter(size_t m) : mtu(m) {}
    uint16_t next_identification() {  // :75-81 wrapping_add
        const uint16_t id = next_id;
        next_id = uint16_t(next_id + 1);
        return id;
    }
    // returns "" and fills `out`, or the reference's Err string
    std::string fragment_packet(const Bytes& packet, std::vector<Bytes>& out) {  // :85-203
        out.clear();
        if (packet.size() <= mtu) {
            out.push_back(packet);
            return "";
        }
        if (packet.size() < 20) return "Invalid IP packet: too small for header";
        const size_t ihl = size_t(packet[0] & 0x0F) * 4;
        if (ihl < 20 || ihl > packet.size()) return "Invalid IP header length";
        if (mtu < ihl) return "MTU too small for fragmentation (need at least 8 bytes of data)";  // `self.mtu - ihl` underflows there
        const size_t max_data = ((mtu - ihl) / 8) * 8;
        if (max_data == 0) return "MTU too small for fragmentation (need at least 8 bytes of data)";
        const uint16_t identification = next_identification();
        const uint8_t* data = packet.data() + 20;  // `&packet[20..]`, whatever ihl says (:116)
        const size_t dlen = packet.size() - 20;
        size_t offset = 0;
        while (offset < dlen) {
            const size_t chunk = std::min(max_data, dlen - offset);
            const bool more = offset + chunk < dlen;
            const uint16_t fo = FragmentationInfo{identification, more, uint16_t(offset / 8)}.to_u16();
            Bytes f(packet.begin(), packet.begin() + 20);
            f[6] = uint8_t(fo >> 8);
            f[7] = uint8_t(fo);
            const uint16_t tl = uint16_t(ihl + chunk);
            f[2] = uint8_t(tl >> 8);
            f[3] = uint8_t(tl);
            f[4] = uint8_t(identification >> 8);
            f[5] = uint8_t(identification);
            f[10] = packet[10];  // checksum zeroed then copied back from the original (:172-178)
            f[11] = packet[11];
            if (ihl > 20) f.insert(f.end(), packet.begin() + 20, packet.begin() + ihl);
            f.insert(f.end(), data + offset, data + offset + chunk);
            out.push_back(std::move(f));
            offset += chunk;
        }
        return "";
    }
};

struct IpReassembler {  // fragmentation.rs:207-401
    using Key = std::pair<uint16_t, std::array<uint8_t, 4>>;
    std::map<Key, std::vector<std::pair<uint16_t, Bytes>>> fragments;
    std::map<Key, bool> last_fragment_seen;
    std::map<Key, Bytes> headers;
    size_t joined = 0;  // fragments in the datagram returned last (for the record; not in the reference)

    // 0 = Ok(None), 1 = Ok(Some(out)), -1 / -2 = the two Err returns
    int process_fragment(const uint8_t* packet, size_t len, Bytes& out) {  // :232-401
        if (len < 20) return -1;
        const size_t ihl = size_t(packet[0] & 0x0F) * 4;
        if (ihl < 20 || ihl > len) return -2;
        const uint16_t flags_offset = uint16_t(packet[6] << 8 | packet[7]);
        const FragmentationInfo fi = FragmentationInfo::from_u16(flags_offset);
        const uint16_t identification = uint16_t(packet[4] << 8 | packet[5]);
        const Key key{identification, {packet[12], packet[13], packet[14], packet[15]}};
        if (!fi.more_fragments && fi.fragment_offset == 0) {  // :264-268
            out.assign(packet, packet + len);
            joined = 1;
            return 1;
        }
        if (!headers.count(key)) headers[key] = Bytes(packet, packet + ihl);
        auto& list = fragments[key];
        list.emplace_back(fi.fragment_offset, Bytes(packet + ihl, packet + len));
        if (!fi.more_fragments) last_fragment_seen[key] = true;
        auto ls = last_fragment_seen.find(key);
        if (ls != last_fragment_seen.end() && ls->second) {
            auto sorted = list;
            std::stable_sort(sorted.begin(), sorted.end(), [](const auto& x, const auto& y) { return x.first < y.first; });
            uint16_t expected = 0;
            bool complete = true;
            for (const auto& [offset, payload] : sorted) {
                if (offset != expected) {
                    complete = false;
                    break;
                }
                expected = uint16_t(offset + uint16_t((payload.size() + 7) / 8));
            }
            if (complete) {
                out = headers[key];
                for (const auto& fp : sorted) out.insert(out.end(), fp.second.begin(), fp.second.end());
                const uint16_t total = uint16_t(out.size());
                if (out.size() >= 4) {
                    out[2] = uint8_t(total >> 8);
                    out[3] = uint8_t(total);
                }
                if (out.size() >= 8) out[6] = out[7] = 0;
                joined = sorted.size();
                fragments.erase(key);
                last_fragment_seen.erase(key);
                headers.erase(key);
                return 1;
            }
        }
        return 0;
    }
};

// dump_to_wav (dump.rs:37-40): (sample * i16::MAX as f32) as i16 — Rust's saturating, truncating cast, NaN -> 0
inline int16_t pcm16_of(float s) {
    const float v = s * 32767.0f;
    if (std::isnan(v)) return 0;
    if (v >= 32767.0f) return 32767;
    if (v <= -32768.0f) return -32768;
    return int16_t(v);  // truncation toward zero
}
// convert_audio_buffer_to_f32, S16 (codec.rs:121-125)
inline float f32_of_pcm16(int16_t v) { return float(v) / 32768.0f; }

}  // namespace orc
