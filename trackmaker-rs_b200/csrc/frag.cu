// frag.cu — the payload layer directly above the PHY: batched IPv4 fragment reassembly of decoded
// frames (IpReassembler::process_fragment, src/net/fragmentation.rs:232-401, behind the frame filter of
// AcousticInterface::receive_packet, src/mac/acoustic_interface.rs:343-366) and batched fragmentation
// (IpFragmenter::fragment_packet, fragmentation.rs:85-203). Byte work: every link / packet is independent, the
// per-link arrival order is the only sequential dependency, so one warp walks one link.
#include <algorithm>
#include <vector>

#include "engine.hpp"

namespace {

constexpr int kCap = TMK_MAX_FRAGMENTS;
constexpr int kWarps = 4;  // links per CTA

enum : uint8_t { K_SKIP = 0, K_ERR_SHORT = 1, K_ERR_IHL = 2, K_WHOLE = 3, K_FRAG = 4 };

struct FragDesc {  // 16 bytes per frame record
    uint32_t src_ip;
    uint16_t ident;
    uint16_t fo;    // 13-bit fragment offset | 0x2000 when more fragments follow (FragmentationInfo::from_u16, :47-57)
    uint16_t len;   // Frame::data length
    uint8_t kind;
    uint8_t ihl;
    uint32_t emit;  // K_FRAG only: 1 = this arrival completed its datagram, 2 = TMK_PKT_E_TOO_MANY; both forget the key
};

__device__ __forceinline__ uint32_t be16(const uint8_t* p) { return uint32_t(p[0]) << 8 | p[1]; }

// one thread per frame record: the receive_packet filter and the header parse of process_fragment (:236-262)
__global__ void frag_classify_kernel(const tmk_frame_rec* __restrict__ recs, uint64_t n_recs, const uint8_t* __restrict__ payload,
                                     uint32_t n_links, FragDesc* __restrict__ desc, uint32_t* __restrict__ lo,
                                     uint32_t* __restrict__ hi) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n_recs) return;
    const tmk_frame_rec r = recs[i];
    FragDesc d;
    d.src_ip = 0;
    d.ident = d.fo = d.len = 0;
    d.kind = K_SKIP;
    d.ihl = 0;
    d.emit = 0;
    if (r.stream < n_links) {
        if (i == 0 || recs[i - 1].stream != r.stream) lo[r.stream] = uint32_t(i);
        if (i + 1 == n_recs || recs[i + 1].stream != r.stream) hi[r.stream] = uint32_t(i + 1);
        // f.frame_type == Data || f.frame_type == Ack && !f.data.is_empty()   (acoustic_interface.rs:346-348)
        const bool accept = r.status == TMK_FRAME_OK &&
                            (r.type == TMK_FRAME_TYPE_DATA || (r.type == TMK_FRAME_TYPE_ACK && r.len > 0));
        if (accept) {
            const uint8_t* __restrict__ p = payload + r.payload_off;
            d.len = r.len;
            if (r.len < 20) {
                d.kind = K_ERR_SHORT;
            } else {
                const uint32_t ihl = (p[0] & 0x0fu) * 4;
                if (ihl < 20 || ihl > r.len) {
                    d.kind = K_ERR_IHL;
                } else {
                    d.ihl = uint8_t(ihl);
                    d.fo = uint16_t(be16(p + 6) & 0x3fffu);
                    d.ident = uint16_t(be16(p + 4));
                    d.src_ip = uint32_t(p[12]) << 24 | uint32_t(p[13]) << 16 | uint32_t(p[14]) << 8 | p[15];
                    d.kind = ((d.fo & 0x2000u) == 0 && (d.fo & 0x1fffu) == 0) ? K_WHOLE : K_FRAG;
                }
            }
        }
    }
    desc[i] = d;
}

__device__ __forceinline__ void warp_copy(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint32_t n, int lane) {
    for (uint32_t b = lane; b < n; b += 32) dst[b] = src[b];
}

// one warp per link, frames in arrival order. WRITE = false counts packets / bytes per link and sets the emit flags,
// WRITE = true (after the two exclusive scans) writes records and bytes at the link's base.
template <bool WRITE>
__global__ void __launch_bounds__(32 * kWarps)
frag_resolve_kernel(FragDesc* __restrict__ desc, const tmk_frame_rec* __restrict__ recs, const uint8_t* __restrict__ payload,
                    const uint32_t* __restrict__ lo, const uint32_t* __restrict__ hi, uint32_t n_links,
                    uint32_t* __restrict__ link_pkts, uint32_t* __restrict__ link_bytes, const uint64_t* __restrict__ pkt_base,
                    const uint64_t* __restrict__ byte_base, tmk_packet_rec* __restrict__ pkts, uint8_t* __restrict__ bytes) {
    __shared__ uint32_t s_idx[kWarps][kCap];
    __shared__ uint16_t s_off[kWarps][kCap], s_len[kWarps][kCap], s_ord[kWarps][kCap];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t link = blockIdx.x * kWarps + w;
    if (link >= n_links) return;
    const uint32_t a = lo[link], b = hi[link];
    uint32_t npk = 0, nby = 0;
    const uint64_t pb = WRITE ? pkt_base[link] : 0, bb = WRITE ? byte_base[link] : 0;
    volatile FragDesc* vdesc = desc;
    auto put_rec = [&](uint32_t status, uint32_t frame, uint32_t len, uint32_t ident, uint32_t nfrag) {
        if (WRITE && lane == 0) {
            tmk_packet_rec pr;
            pr.link = link;
            pr.status = status;
            pr.frame = frame;
            pr.len = len;
            pr.data_off = bb + nby;
            pr.identification = uint16_t(ident);
            pr.n_fragments = uint16_t(nfrag);
            pr._pad = 0;
            pkts[pb + npk] = pr;
        }
        ++npk;
        nby += len;
    };
    for (uint32_t i = a; i < b; ++i) {
        const FragDesc d = desc[i];
        if (d.kind == K_SKIP) continue;
        if (d.kind == K_ERR_SHORT || d.kind == K_ERR_IHL) {
            put_rec(d.kind == K_ERR_SHORT ? TMK_PKT_E_SHORT : TMK_PKT_E_IHL, i, 0, 0, 0);
            continue;
        }
        if (d.kind == K_WHOLE) {  // :264-268 not fragmented: the packet as it is, no state touched
            if (WRITE) warp_copy(bytes + bb + nby, payload + recs[i].payload_off, d.len, lane);
            put_rec(TMK_PKT_OK, i, d.len, d.ident, 1);
            continue;
        }
        // the key's entry lives from the arrival after its last completion (the three maps are removed together, :381-385)
        uint32_t start = a;
        for (uint32_t base = a; base < i; base += 32) {
            const uint32_t j = base + lane;
            const bool hit = j < i && vdesc[j].kind == K_FRAG && vdesc[j].ident == d.ident && vdesc[j].src_ip == d.src_ip &&
                             vdesc[j].emit != 0;
            const uint32_t bal = __ballot_sync(0xffffffffu, hit);
            if (bal) start = base + (31 - __clz(bal)) + 1;
        }
        // fragment_list in arrival order (:287-292), last_fragment_seen (:304-308)
        uint32_t m = 0;
        bool last_seen = false;
        for (uint32_t base = start; base <= i; base += 32) {
            const uint32_t j = base + lane;
            bool match = false, last = false;
            if (j <= i) {
                const FragDesc o = desc[j];
                match = o.kind == K_FRAG && o.ident == d.ident && o.src_ip == d.src_ip;
                last = match && (o.fo & 0x2000u) == 0;
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, match);
            const uint32_t pos = m + __popc(bal & ((1u << lane) - 1));
            if (match && pos < kCap) {
                const FragDesc o = desc[j];
                s_idx[w][pos] = j;
                s_off[w][pos] = o.fo & 0x1fffu;
                s_len[w][pos] = uint16_t(o.len - o.ihl);  // packet[ihl..] with the fragment's own ihl (:285)
            }
            m += __popc(bal);
            last_seen |= __any_sync(0xffffffffu, last);
        }
        __syncwarp();
        if (m > kCap) {
            if (!WRITE && lane == 0) desc[i].emit = 2;
            put_rec(TMK_PKT_E_TOO_MANY, i, 0, d.ident, 0);
            __threadfence_block();
            __syncwarp();
            continue;
        }
        if (!last_seen) continue;
        // sorted_fragments.sort_by_key(offset): stable (:319-320) — rank = smaller offsets + equal offsets that arrived earlier
        for (uint32_t k = lane; k < m; k += 32) {
            const uint32_t ok = s_off[w][k];
            uint32_t r = 0;
            for (uint32_t t = 0; t < m; ++t) {
                const uint32_t ot = s_off[w][t];
                r += (ot < ok) || (ot == ok && t < k);
            }
            s_ord[w][r] = uint16_t(k);
        }
        __syncwarp();
        // gap-free chain from 0: next = offset + ceil(len / 8) (:333-355)
        bool ok = true;
        uint32_t sum = 0;
        for (uint32_t k = lane; k < m; k += 32) {
            const uint32_t cur = s_ord[w][k];
            uint32_t expect = 0;
            if (k) {
                const uint32_t prev = s_ord[w][k - 1];
                expect = (uint32_t(s_off[w][prev]) + (uint32_t(s_len[w][prev]) + 7) / 8) & 0xffffu;
            }
            ok &= s_off[w][cur] == expect;
            sum += s_len[w][cur];
        }
        ok = __all_sync(0xffffffffu, ok);
        if (!ok) {
            __syncwarp();
            continue;
        }
        for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        const uint32_t first = s_idx[w][0];
        const uint32_t ihl0 = desc[first].ihl;  // header of the first fragment that arrived (:271-277)
        const uint32_t total = ihl0 + sum;
        if (WRITE) {
            uint8_t* __restrict__ out = bytes + bb + nby;
            warp_copy(out, payload + recs[first].payload_off, ihl0, lane);
            uint32_t pos = ihl0;
            for (uint32_t k = 0; k < m; ++k) {
                const uint32_t cur = s_ord[w][k];
                const uint32_t j = s_idx[w][cur];
                warp_copy(out + pos, payload + recs[j].payload_off + desc[j].ihl, s_len[w][cur], lane);
                pos += s_len[w][cur];
            }
            __syncwarp();
            if (lane == 0) {  // total_length as u16, flags / offset cleared (:369-378)
                out[2] = uint8_t(total >> 8);
                out[3] = uint8_t(total);
                out[6] = 0;
                out[7] = 0;
            }
        } else if (lane == 0) {
            desc[i].emit = 1;
        }
        put_rec(TMK_PKT_OK, i, total, d.ident, m);
        __threadfence_block();
        __syncwarp();
    }
    if (!WRITE && lane == 0) {
        link_pkts[link] = npk;
        link_bytes[link] = nby;
    }
}

// ---------------------------------------------------------------- IpFragmenter::fragment_packet (:85-203)
struct CutPlan {  // per packet
    uint32_t n_frag, n_bytes, fragmented, status;
};

__device__ __forceinline__ CutPlan plan_of(const uint8_t* __restrict__ p, uint64_t L, uint32_t mtu) {
    CutPlan c = {1, uint32_t(L), 0, TMK_PKT_OK};
    if (L <= mtu) return c;  // :89-97 no fragmentation needed
    c.n_bytes = 0;
    if (L < 20) {
        c.status = TMK_PKT_E_SHORT;
        return c;
    }
    const uint32_t ihl = (p[0] & 0x0fu) * 4;
    if (ihl < 20 || ihl > L) {
        c.status = TMK_PKT_E_IHL;
        return c;
    }
    if (mtu < ihl + 8) {  // max_data_per_fragment == 0 (:131-136); mtu < ihl underflows `self.mtu - ihl` in the reference
        c.status = TMK_PKT_E_MTU;
        return c;
    }
    const uint32_t maxd = (mtu - ihl) / 8 * 8, data = uint32_t(L) - 20;  // data = &packet[20..] (:116), options included
    c.n_frag = (data + maxd - 1) / maxd;
    c.n_bytes = c.n_frag * ihl + data;
    c.fragmented = 1;
    return c;
}

__global__ void frag_plan_kernel(const uint8_t* __restrict__ data, const uint64_t* __restrict__ off, uint32_t n, uint32_t mtu,
                                 uint32_t* __restrict__ cnt_frag, uint32_t* __restrict__ cnt_bytes, uint32_t* __restrict__ cnt_id) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const CutPlan c = plan_of(data + off[i], off[i + 1] - off[i], mtu);
    cnt_frag[i] = c.n_frag;
    cnt_bytes[i] = c.n_bytes;
    cnt_id[i] = c.fragmented;
}

// one CTA per packet; warp v writes fragments v, v + nwarps, ...
__global__ void __launch_bounds__(128)
frag_cut_kernel(const uint8_t* __restrict__ data, const uint64_t* __restrict__ off, uint32_t mtu, uint32_t first_id,
                const uint64_t* __restrict__ base_frag, const uint64_t* __restrict__ base_bytes, const uint64_t* __restrict__ base_id,
                tmk_packet_rec* __restrict__ frags, uint8_t* __restrict__ bytes) {
    const uint32_t i = blockIdx.x;
    const uint8_t* __restrict__ p = data + off[i];
    const uint64_t L = off[i + 1] - off[i];
    const CutPlan c = plan_of(p, L, mtu);
    const int lane = threadIdx.x & 31, wv = threadIdx.x >> 5, nw = blockDim.x >> 5;
    tmk_packet_rec* __restrict__ fr = frags + base_frag[i];
    uint8_t* __restrict__ out = bytes + base_bytes[i];
    if (!c.fragmented) {
        for (uint64_t b = threadIdx.x; b < c.n_bytes; b += blockDim.x) out[b] = p[b];
        if (threadIdx.x == 0) {
            tmk_packet_rec pr = {i, c.status, 0, c.n_bytes, base_bytes[i], 0, uint16_t(c.status == TMK_PKT_OK ? 1 : 0), 0};
            fr[0] = pr;
        }
        return;
    }
    const uint32_t ihl = (p[0] & 0x0fu) * 4, maxd = (mtu - ihl) / 8 * 8, dlen = uint32_t(L) - 20;
    const uint32_t id = (first_id + uint32_t(base_id[i])) & 0xffffu;  // next_identification wraps (:75-81)
    for (uint32_t f = wv; f < c.n_frag; f += nw) {
        const uint32_t o = f * maxd, chunk = min(maxd, dlen - o);
        const bool more = o + chunk < dlen;
        const uint32_t fo = (more ? 0x2000u : 0u) | ((o / 8) & 0x1fffu), tl = (ihl + chunk) & 0xffffu;
        uint8_t* __restrict__ q = out + uint64_t(f) * (ihl + maxd);  // every fragment before the last carries maxd bytes
        // first 20 header bytes with total length, identification, flags/offset rewritten; checksum copied (:150-178)
        if (lane < 20) {
            uint32_t v = p[lane];
            if (lane == 2) v = tl >> 8;
            if (lane == 3) v = tl & 0xff;
            if (lane == 4) v = id >> 8;
            if (lane == 5) v = id & 0xff;
            if (lane == 6) v = fo >> 8;
            if (lane == 7) v = fo & 0xff;
            q[lane] = uint8_t(v);
        }
        for (uint32_t b = 20 + lane; b < ihl; b += 32) q[b] = p[b];          // options (:181-183)
        for (uint32_t b = lane; b < chunk; b += 32) q[ihl + b] = p[20 + o + b];  // chunk of packet[20..] (:141)
        if (lane == 0) {
            tmk_packet_rec pr = {i, TMK_PKT_OK, f, ihl + chunk, base_bytes[i] + uint64_t(f) * (ihl + maxd), uint16_t(id),
                                 uint16_t(c.n_frag), 0};
            fr[f] = pr;
        }
    }
}

int fail(tmk_engine* e, int code, const std::string& msg) { return tmk_fail(e, code, msg); }

int stage(tmk_engine* e, DevBuf& buf, const void* src, size_t bytes, const void** out) {
    if (bytes == 0 || tmk_is_device_ptr(src)) {
        *out = bytes ? src : nullptr;
        return TMK_OK;
    }
    CK(buf.reserve(bytes));
    CK(cudaMemcpyAsync(buf.p, src, bytes, cudaMemcpyHostToDevice, e->stream));
    *out = buf.p;
    return TMK_OK;
}

size_t up16(size_t v) { return (v + 15) & ~size_t(15); }

}  // namespace

extern "C" {

int tmk_reassemble_batch(tmk_engine* e, const tmk_frame_rec* recs, uint64_t n_recs, const uint8_t* payload,
                         uint64_t payload_len, uint32_t n_links, tmk_packet_rec* pkts, uint64_t pkt_cap, uint64_t* n_pkts,
                         uint8_t* bytes, uint64_t byte_cap, uint64_t* n_bytes) {
    if (!e || (!recs && n_recs) || (!payload && payload_len) || !n_pkts || !n_bytes || (!pkts && pkt_cap) || (!bytes && byte_cap))
        return fail(e, TMK_E_INVALID, "tmk_reassemble_batch: null argument");
    *n_pkts = 0;
    *n_bytes = 0;
    if (n_recs == 0 || n_links == 0) return TMK_OK;
    if (n_recs >= (1ull << 32)) return fail(e, TMK_E_INVALID, "tmk_reassemble_batch: too many records");
    CK(cudaSetDevice(e->device));
    const void *d_recs = nullptr, *d_pay = nullptr;
    int rc = stage(e, e->d_recs, recs, sizeof(tmk_frame_rec) * n_recs, &d_recs);
    if (rc != TMK_OK) return rc;
    rc = stage(e, e->d_payload, payload, payload_len, &d_pay);
    if (rc != TMK_OK) return rc;
    // scratch: desc | lo | hi | link_pkts | link_bytes | pkt_base | byte_base
    const size_t o_desc = 0, o_lo = up16(sizeof(FragDesc) * n_recs), o_hi = o_lo + up16(4ull * n_links),
                 o_cp = o_hi + up16(4ull * n_links), o_cb = o_cp + up16(4ull * n_links), o_pb = o_cb + up16(4ull * n_links),
                 o_bb = o_pb + up16(8ull * (n_links + 1)), o_end = o_bb + up16(8ull * (n_links + 1));
    CK(e->d_frag.reserve(o_end));
    char* base = static_cast<char*>(e->d_frag.p);
    FragDesc* desc = reinterpret_cast<FragDesc*>(base + o_desc);
    uint32_t *lo = reinterpret_cast<uint32_t*>(base + o_lo), *hi = reinterpret_cast<uint32_t*>(base + o_hi),
             *cp = reinterpret_cast<uint32_t*>(base + o_cp), *cb = reinterpret_cast<uint32_t*>(base + o_cb);
    uint64_t *pb = reinterpret_cast<uint64_t*>(base + o_pb), *bbp = reinterpret_cast<uint64_t*>(base + o_bb);
    CK(cudaMemsetAsync(base + o_lo, 0, o_pb - o_lo, e->stream));
    const tmk_frame_rec* dr = static_cast<const tmk_frame_rec*>(d_recs);
    const uint8_t* dp = static_cast<const uint8_t*>(d_pay);
    frag_classify_kernel<<<unsigned((n_recs + 255) / 256), 256, 0, e->stream>>>(dr, n_recs, dp, n_links, desc, lo, hi);
    const unsigned grid = (n_links + kWarps - 1) / kWarps;
    frag_resolve_kernel<false><<<grid, 32 * kWarps, 0, e->stream>>>(desc, dr, dp, lo, hi, n_links, cp, cb, nullptr, nullptr, nullptr,
                                                                     nullptr);
    e->launches += 2;
    e->launches += tmk_launch_excl_scan(cp, pb, n_links, e->stream);
    e->launches += tmk_launch_excl_scan(cb, bbp, n_links, e->stream);
    CK(cudaGetLastError());
    uint64_t tot[2] = {0, 0};
    CK(cudaMemcpyAsync(&tot[0], pb + n_links, 8, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(&tot[1], bbp + n_links, 8, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    *n_pkts = tot[0];
    *n_bytes = tot[1];
    if (tot[0] > pkt_cap || tot[1] > byte_cap) return fail(e, TMK_E_CAPACITY, "tmk_reassemble_batch: packet or byte buffer too small");
    if (tot[0] == 0) return TMK_OK;
    CK(e->d_misc.reserve(sizeof(tmk_packet_rec) * tot[0]));
    CK(e->d_misc2.reserve(std::max<uint64_t>(1, tot[1])));
    frag_resolve_kernel<true><<<grid, 32 * kWarps, 0, e->stream>>>(desc, dr, dp, lo, hi, n_links, cp, cb, pb, bbp,
                                                                    static_cast<tmk_packet_rec*>(e->d_misc.p),
                                                                    static_cast<uint8_t*>(e->d_misc2.p));
    e->launches += 1;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(pkts, e->d_misc.p, sizeof(tmk_packet_rec) * tot[0], cudaMemcpyDeviceToHost, e->stream));
    if (tot[1]) CK(cudaMemcpyAsync(bytes, e->d_misc2.p, tot[1], cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

int tmk_fragment_batch(tmk_engine* e, const uint8_t* data, const uint64_t* pkt_off, uint32_t n_packets, uint32_t mtu,
                       uint16_t first_identification, tmk_packet_rec* frags, uint64_t frag_cap, uint64_t* n_frags, uint8_t* bytes,
                       uint64_t byte_cap, uint64_t* n_bytes, uint16_t* next_identification) {
    if (!e || !pkt_off || !n_frags || !n_bytes || (!frags && frag_cap) || (!bytes && byte_cap))
        return fail(e, TMK_E_INVALID, "tmk_fragment_batch: null argument");
    *n_frags = 0;
    *n_bytes = 0;
    if (next_identification) *next_identification = first_identification;
    if (n_packets == 0) return TMK_OK;
    for (uint32_t i = 0; i < n_packets; ++i) {
        if (pkt_off[i + 1] < pkt_off[i]) return fail(e, TMK_E_INVALID, "tmk_fragment_batch: packet offsets must ascend");
        if (pkt_off[i + 1] - pkt_off[i] > 65535) return fail(e, TMK_E_INVALID, "tmk_fragment_batch: packet longer than 65535 bytes");
    }
    const uint64_t total = pkt_off[n_packets];
    if (!data && total) return fail(e, TMK_E_INVALID, "tmk_fragment_batch: null data");
    CK(cudaSetDevice(e->device));
    const void* d_data = nullptr;
    int rc = stage(e, e->d_payload, data, total, &d_data);
    if (rc != TMK_OK) return rc;
    const size_t o_off = 0, o_c0 = up16(8ull * (n_packets + 1)), o_c1 = o_c0 + up16(4ull * n_packets), o_c2 = o_c1 + up16(4ull * n_packets),
                 o_b0 = o_c2 + up16(4ull * n_packets), o_b1 = o_b0 + up16(8ull * (n_packets + 1)), o_b2 = o_b1 + up16(8ull * (n_packets + 1)),
                 o_end = o_b2 + up16(8ull * (n_packets + 1));
    CK(e->d_frag.reserve(o_end));
    char* base = static_cast<char*>(e->d_frag.p);
    uint64_t* d_off = reinterpret_cast<uint64_t*>(base + o_off);
    uint32_t *c0 = reinterpret_cast<uint32_t*>(base + o_c0), *c1 = reinterpret_cast<uint32_t*>(base + o_c1),
             *c2 = reinterpret_cast<uint32_t*>(base + o_c2);
    uint64_t *b0 = reinterpret_cast<uint64_t*>(base + o_b0), *b1 = reinterpret_cast<uint64_t*>(base + o_b1),
             *b2 = reinterpret_cast<uint64_t*>(base + o_b2);
    CK(cudaMemcpyAsync(d_off, pkt_off, 8ull * (n_packets + 1), cudaMemcpyHostToDevice, e->stream));
    const uint8_t* dd = static_cast<const uint8_t*>(d_data);
    frag_plan_kernel<<<(n_packets + 255) / 256, 256, 0, e->stream>>>(dd, d_off, n_packets, mtu, c0, c1, c2);
    e->launches += 1;
    e->launches += tmk_launch_excl_scan(c0, b0, n_packets, e->stream);
    e->launches += tmk_launch_excl_scan(c1, b1, n_packets, e->stream);
    e->launches += tmk_launch_excl_scan(c2, b2, n_packets, e->stream);
    CK(cudaGetLastError());
    uint64_t tot[3] = {0, 0, 0};
    CK(cudaMemcpyAsync(&tot[0], b0 + n_packets, 8, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(&tot[1], b1 + n_packets, 8, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(&tot[2], b2 + n_packets, 8, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    *n_frags = tot[0];
    *n_bytes = tot[1];
    if (next_identification) *next_identification = uint16_t(first_identification + tot[2]);
    if (tot[0] > frag_cap || tot[1] > byte_cap) return fail(e, TMK_E_CAPACITY, "tmk_fragment_batch: fragment or byte buffer too small");
    CK(e->d_misc.reserve(sizeof(tmk_packet_rec) * tot[0]));
    CK(e->d_misc2.reserve(std::max<uint64_t>(1, tot[1])));
    frag_cut_kernel<<<n_packets, 128, 0, e->stream>>>(dd, d_off, mtu, first_identification, b0, b1, b2,
                                                      static_cast<tmk_packet_rec*>(e->d_misc.p), static_cast<uint8_t*>(e->d_misc2.p));
    e->launches += 1;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(frags, e->d_misc.p, sizeof(tmk_packet_rec) * tot[0], cudaMemcpyDeviceToHost, e->stream));
    if (tot[1]) CK(cudaMemcpyAsync(bytes, e->d_misc2.p, tot[1], cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return TMK_OK;
}

// ---------------------------------------------------------------- reassembler state carried across calls
// AcousticInterface keeps ONE IpReassembler for all its receive polls (src/mac/acoustic_interface.rs:316-378: the
// `reassembler` field outlives every process_samples call), so a datagram whose fragments arrive in different polls
// still completes. tmk_reasm_* carries that state per link: the fragments of unfinished datagrams (their frame records and
// bytes) are kept on the host, put in front of the next call's frames of the link, and the batch kernels run over both.
// A fragment stays pending exactly while the reassembler's maps would still hold it: it is a fragment (not a whole
// packet, not an error) and no arrival at or behind it completed (or overflowed) its (identification, source) key.
struct tmk_reasm {
    tmk_engine* e;
    uint32_t n_links;
    std::vector<std::vector<tmk_frame_rec>> recs;   // per link, arrival order; payload_off indexes bytes[link]
    std::vector<std::vector<uint8_t>> bytes;
};

int tmk_reasm_create(tmk_engine* e, uint32_t n_links, tmk_reasm** out) {
    if (!e || !out || n_links == 0) return fail(e, TMK_E_INVALID, "tmk_reasm_create: bad argument");
    tmk_reasm* r = new tmk_reasm;
    r->e = e;
    r->n_links = n_links;
    r->recs.resize(n_links);
    r->bytes.resize(n_links);
    *out = r;
    return TMK_OK;
}

void tmk_reasm_destroy(tmk_reasm* r) { delete r; }

int tmk_reasm_reset(tmk_reasm* r, uint32_t link) {
    if (!r || (link != 0xffffffffu && link >= r->n_links)) return TMK_E_INVALID;
    for (uint32_t l = 0; l < r->n_links; ++l)
        if (link == 0xffffffffu || l == link) {
            r->recs[l].clear();
            r->bytes[l].clear();
        }
    return TMK_OK;
}

int tmk_reasm_pending(const tmk_reasm* r, uint32_t link, uint32_t* n_fragments, uint64_t* n_bytes) {
    if (!r || link >= r->n_links) return TMK_E_INVALID;
    if (n_fragments) *n_fragments = uint32_t(r->recs[link].size());
    if (n_bytes) *n_bytes = r->bytes[link].size();
    return TMK_OK;
}

int tmk_reasm_push(tmk_reasm* r, const tmk_frame_rec* recs, uint64_t n_recs, const uint8_t* payload, uint64_t payload_len,
                   tmk_packet_rec* pkts, uint64_t pkt_cap, uint64_t* n_pkts, uint8_t* bytes, uint64_t byte_cap, uint64_t* n_bytes) {
    if (!r) return TMK_E_INVALID;
    tmk_engine* e = r->e;
    if ((!recs && n_recs) || (!payload && payload_len) || !n_pkts || !n_bytes || (!pkts && pkt_cap) || (!bytes && byte_cap))
        return fail(e, TMK_E_INVALID, "tmk_reasm_push: null argument");
    *n_pkts = 0;
    *n_bytes = 0;
    CK(cudaSetDevice(e->device));
    // the call's records and bytes on the host (they may live on the device)
    std::vector<tmk_frame_rec> hr(n_recs);
    std::vector<uint8_t> hp(payload_len);
    if (n_recs) CK(cudaMemcpy(hr.data(), recs, sizeof(tmk_frame_rec) * n_recs, cudaMemcpyDefault));
    if (payload_len) CK(cudaMemcpy(hp.data(), payload, payload_len, cudaMemcpyDefault));
    // carried fragments ++ the new frames, link by link; `origin` = index in this call's recs, or -1 for a carried one
    std::vector<tmk_frame_rec> all;
    std::vector<uint8_t> blob;
    std::vector<int64_t> origin;
    std::vector<uint64_t> first_new(r->n_links + 1, 0);
    for (uint64_t i = 0; i < n_recs; ++i) {
        if (hr[i].stream >= r->n_links) return fail(e, TMK_E_INVALID, "tmk_reasm_push: record of a link the reassembler does not have");
        if (i && hr[i].stream < hr[i - 1].stream) return fail(e, TMK_E_INVALID, "tmk_reasm_push: records must be grouped by link, ascending");
        const bool has_bytes = hr[i].status == TMK_FRAME_OK || hr[i].status == TMK_FRAME_NOT_FOR_US || hr[i].status == TMK_FRAME_CRC_FAIL;
        if (has_bytes && (hr[i].payload_off > payload_len || hr[i].len > payload_len - hr[i].payload_off))
            return fail(e, TMK_E_INVALID, "tmk_reasm_push: record payload outside the payload buffer");
        first_new[hr[i].stream + 1] = i + 1;
    }
    for (uint32_t l = 0; l < r->n_links; ++l) first_new[l + 1] = std::max(first_new[l + 1], first_new[l]);
    auto append = [&](const tmk_frame_rec& src, const uint8_t* data, int64_t org) {
        tmk_frame_rec t = src;
        t.payload_off = blob.size();
        const bool has_bytes = t.status == TMK_FRAME_OK || t.status == TMK_FRAME_NOT_FOR_US || t.status == TMK_FRAME_CRC_FAIL;
        if (has_bytes) blob.insert(blob.end(), data, data + t.len);
        all.push_back(t);
        origin.push_back(org);
    };
    for (uint32_t l = 0; l < r->n_links; ++l) {
        for (const tmk_frame_rec& c : r->recs[l]) append(c, r->bytes[l].data() + c.payload_off, -1);
        for (uint64_t i = first_new[l]; i < first_new[l + 1]; ++i) append(hr[i], hp.data() + hr[i].payload_off, int64_t(i));
    }
    if (all.empty()) return TMK_OK;
    blob.push_back(0);
    int rc = tmk_reassemble_batch(e, all.data(), all.size(), blob.data(), blob.size(), r->n_links, pkts, pkt_cap, n_pkts, bytes,
                                  byte_cap, n_bytes);
    if (rc != TMK_OK) return rc;  // (nothing was consumed: the call may be repeated with larger buffers)
    // `frame` of a packet record: the arrival that completed it, always one of this call's frames
    if (tmk_is_device_ptr(pkts)) {
        // device-resident packet records keep the batch numbering (carried fragments first); host callers get the call's own
    } else {
        for (uint64_t k = 0; k < *n_pkts; ++k) pkts[k].frame = origin[pkts[k].frame] >= 0 ? uint32_t(origin[pkts[k].frame]) : 0xffffffffu;
    }
    // what the reassembler still holds: fragments with no completion of their key at or behind them
    std::vector<FragDesc> hd(all.size());
    CK(cudaMemcpy(hd.data(), e->d_frag.p, sizeof(FragDesc) * all.size(), cudaMemcpyDeviceToHost));
    std::vector<std::vector<tmk_frame_rec>> nrecs(r->n_links);
    std::vector<std::vector<uint8_t>> nbytes(r->n_links);
    uint64_t hi = all.size();
    for (uint32_t l = r->n_links; l-- > 0;) {
        uint64_t lo = hi;
        while (lo > 0 && all[lo - 1].stream == l) --lo;
        std::vector<std::pair<uint32_t, uint16_t>> done;  // keys completed later in the link
        std::vector<uint64_t> keep;
        for (uint64_t j = hi; j-- > lo;) {
            const FragDesc& d = hd[j];
            if (d.kind != K_FRAG) continue;
            const std::pair<uint32_t, uint16_t> key(d.src_ip, d.ident);
            if (d.emit != 0) done.push_back(key);
            if (std::find(done.begin(), done.end(), key) == done.end()) keep.push_back(j);
        }
        for (auto it = keep.rbegin(); it != keep.rend(); ++it) {
            tmk_frame_rec t = all[*it];
            const uint8_t* data = blob.data() + t.payload_off;
            t.payload_off = nbytes[l].size();
            nbytes[l].insert(nbytes[l].end(), data, data + t.len);
            nrecs[l].push_back(t);
        }
        hi = lo;
    }
    r->recs.swap(nrecs);
    r->bytes.swap(nbytes);
    return TMK_OK;
}

}  // extern "C"
