// workload.cpp — the synthetic workload of SURVEY.md §8(d), written a second time.  TEST INFRASTRUCTURE
// ONLY (part of libsfc_oracle.so): bench.py's `--impl reference` / cpu_baseline legs draw their inputs from
// here so that they never load the product library, and tests/test_oracle.py holds this file and the
// product's generator (dpu-operator_b200/csrc/gen.h) to identical bytes — two writings of one
// specification, so an error in either shows up as a difference.
//
// Specification (integer only).  R(seed, k) = the k-th output of splitmix64 started at `seed`:
// finalise(seed + (k + 1) * 0x9E3779B97F4A7C15).  Flow i draws R(seed_rules, 4i .. 4i+3), packet i draws
// R(seed_pkts, 4i .. 4i+3), prefix j draws R(seed_lpm [^ 0x66.. for IPv6], 2j .. 2j+1), ACL rule j draws
// R(seed_acl, 4j .. 4j+3).  Which bits mean what is spelled out beside each use below.
#include <cstdint>
#include <cstring>

#include "../include/b200sfc.h"

namespace {

uint64_t finalise(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
uint64_t R(uint64_t seed, uint64_t k) { return finalise(seed + (k + 1) * 0x9E3779B97F4A7C15ull); }
uint32_t hi32(uint64_t v) { return (uint32_t)(v >> 32); }
uint32_t lo32(uint64_t v) { return (uint32_t)v; }
uint32_t top_bits(int n) { return n <= 0 ? 0u : n >= 32 ? 0xFFFFFFFFu : 0xFFFFFFFFu << (32 - n); }

struct Bytes {  // big-endian writer over a zeroed 64-byte record
    uint8_t* p;
    void u8(int at, uint32_t v) { p[at] = (uint8_t)v; }
    void u16(int at, uint32_t v) {
        p[at] = (uint8_t)(v >> 8);
        p[at + 1] = (uint8_t)v;
    }
    void u32(int at, uint32_t v) {
        u16(at, v >> 16);
        u16(at + 2, v & 0xFFFFu);
    }
};

// VF MAC pattern of the reference: 00:<vsi>:00:00:03:14 (vendored ipuplugin/bridgeport.go:99-100), with the
// index spread over bytes 1..3
void mac_of(uint32_t i, uint8_t* m) {
    m[0] = 0;
    m[1] = (uint8_t)i;
    m[2] = (uint8_t)(i >> 8);
    m[3] = (uint8_t)(i >> 16);
    m[4] = 0x03;
    m[5] = 0x14;
}

// 802.1Q: ten bits of r (bits 1..10) scaled to 0..999 decide, bits 12..23 pick a VID in 1..4094
uint32_t vlan_of(uint64_t r, uint32_t permille) {
    uint32_t scaled = (((uint32_t)(r >> 1) & 1023u) * 1000u) >> 10;
    if (scaled >= permille) return 0;
    return 1u + ((uint32_t)(r >> 12) & 0xFFFu) % 4094u;
}

struct Tuple {
    uint32_t src[4] = {0, 0, 0, 0}, dst[4] = {0, 0, 0, 0};
    uint32_t sport = 0, dport = 0, vlan = 0, proto = 0, ver = 4, action = 0, out_port = 0;
};

Tuple flow_of(const sfc_gen_spec& s, uint32_t i) {
    const uint64_t a = R(s.seed_rules, 4ull * i), b = R(s.seed_rules, 4ull * i + 1), c = R(s.seed_rules, 4ull * i + 2),
                   d = R(s.seed_rules, 4ull * i + 3);
    Tuple t;
    t.ver = hi32(a) % 1000u < s.v6_permille ? 6 : 4;
    t.proto = (a & 1) ? 17 : 6;
    t.vlan = vlan_of(a, s.vlan_permille);
    t.sport = lo32(b) & 0xFFFFu;
    t.dport = (lo32(b) >> 16) & 0xFFFFu;
    const bool drop = (hi32(b) & 0xFFu) < 13u;  // about 5 % of the flows drop
    t.action = drop ? SFC_ACT_DROP : SFC_ACT_FORWARD;
    t.out_port = drop ? 0 : ((uint32_t)(b >> 40) & 0xFFFFu) % (s.n_out_ports ? s.n_out_ports : 1u);
    t.src[0] = hi32(c);
    if (t.ver == 4) {
        t.dst[0] = lo32(c);
    } else {
        const uint64_t e = finalise(c ^ 0x6A09E667F3BCC909ull), f = finalise(d ^ 0xBB67AE8584CAA73Bull);
        t.src[1] = lo32(c);
        t.src[2] = hi32(d);
        t.src[3] = lo32(d);
        t.dst[0] = hi32(e);
        t.dst[1] = lo32(e);
        t.dst[2] = hi32(f);
        t.dst[3] = lo32(f);
    }
    return t;
}

// cumulative per-mille histograms of prefix lengths: IPv4 /8../32 with the peak at /24, IPv6 /32../64 peak /48
int len4_of(uint32_t u) {
    static const uint16_t cum[25] = {1,   2,   3,   4,   6,   9,   14,  19,  59,  74,  94,  129, 174,
                                     219, 309, 379, 939, 944, 949, 954, 962, 972, 984, 985, 1000};
    int k = 0;
    while (k < 25 && u >= cum[k]) ++k;
    return k < 25 ? 8 + k : 32;
}
int len6_of(uint32_t u) {
    static const uint16_t cum[9] = {100, 130, 190, 250, 750, 770, 870, 900, 1000};
    int k = 0;
    while (k < 9 && u >= cum[k]) ++k;
    return k < 9 ? 32 + 4 * k : 64;
}
void prefix_of(const sfc_gen_spec& s, int af, uint32_t j, uint32_t* w, int& len, uint32_t& port) {
    const uint64_t seed = af == 6 ? s.seed_lpm ^ 0x6666666666666666ull : s.seed_lpm;
    const uint64_t r = R(seed, 2ull * j), q = R(seed, 2ull * j + 1);
    port = (lo32(q) & 0xFFFFu) % (s.n_out_ports ? s.n_out_ports : 1u);
    w[0] = w[1] = w[2] = w[3] = 0;
    if (af == 4) {
        len = len4_of(lo32(r) % 1000u);
        w[0] = hi32(r) & top_bits(len);
    } else {
        len = len6_of(lo32(r) % 1000u);
        w[0] = (0x20000000u | (hi32(r) & 0x1FFFFFFFu)) & top_bits(len);  // inside 2000::/3
        w[1] = hi32(q) & top_bits(len - 32);
    }
}

// Which flow a replayed packet belongs to: uniform, or — zipf — a uniformly chosen octave [2^k - 1, 2^(k+1) - 1)
// and a uniform rank inside it, so every octave carries the same share of the traffic
uint32_t flow_pick(const sfc_gen_spec& s, uint64_t p) {
    if (!s.zipf) return (uint32_t)(p % s.n_flows);
    uint32_t top = 0;
    while ((2u << top) <= s.n_flows && top < 31) ++top;  // floor(log2(n_flows))
    const uint32_t k = (uint32_t)(p >> 58) % (top + 1u);
    const uint64_t within = (p & ((1ull << 58) - 1ull)) & ((1ull << k) - 1ull);
    return (uint32_t)((((1ull << k) - 1ull) + within) % s.n_flows);
}

void packet_of(const sfc_gen_spec& s, uint64_t i, uint8_t* rec, uint16_t& in_port) {
    const uint64_t p0 = R(s.seed_pkts, 4 * i), p1 = R(s.seed_pkts, 4 * i + 1), p2 = R(s.seed_pkts, 4 * i + 2),
                   p3 = R(s.seed_pkts, 4 * i + 3);
    memset(rec, 0, 64);
    Bytes o{rec};
    in_port = (uint16_t)(s.port_base + (lo32(p3) & 0xFFFFu) % (s.n_ports ? s.n_ports : 1u));
    mac_of(((uint32_t)(p3 >> 16) & 0xFFFFFFu) % (s.n_macs ? s.n_macs : 1u), rec);
    o.u8(6, 0x02);  // source MAC: locally administered, some varying bytes
    o.u8(9, (uint32_t)(p3 >> 40));
    o.u8(10, (uint32_t)(p3 >> 48));
    o.u8(11, (uint32_t)(p2 >> 8));

    if (hi32(p0) % 1000u < s.nonip_permille) {  // an ARP request, optionally tagged (PCP 1)
        const uint32_t vlan = vlan_of(p1, s.vlan_permille);
        int at = 12;
        if (vlan) {
            o.u16(12, 0x8100);
            o.u16(14, vlan | 0x2000u);
            at = 16;
        }
        o.u16(at, 0x0806);
        o.u32(at + 2, 0x00010800u);
        o.u32(at + 6, 0x06040001u);
        o.u32(at + 10, hi32(p1));
        o.u32(at + 14, lo32(p2));
        o.u32(at + 18, hi32(p2));
        return;
    }
    Tuple t;
    if (s.n_flows > 0 && lo32(p0) % 1000u < s.hit_permille) {
        t = flow_of(s, flow_pick(s, p1));
    } else {  // a packet of no flow: random tuple, destination optionally inside a prefix of the LPM set
        const uint64_t q = finalise(p1);
        static const uint8_t protos[4] = {6, 17, 1, 132};
        t.ver = hi32(p1) % 1000u < s.v6_permille ? 6 : 4;
        t.proto = protos[p1 & 3];
        t.vlan = vlan_of(p1 >> 1, s.vlan_permille);
        t.sport = lo32(q) & 0xFFFFu;
        t.dport = (lo32(q) >> 16) & 0xFFFFu;
        t.src[0] = hi32(p2);
        t.dst[0] = lo32(p2);
        if (t.ver == 6) {
            const uint64_t a = finalise(p2), b = finalise(p2 ^ 0x3C6EF372FE94F82Bull), d = finalise(a ^ b);
            t.src[1] = lo32(a);
            t.src[2] = hi32(a);
            t.src[3] = lo32(b);
            t.dst[1] = hi32(b);
            t.dst[2] = lo32(d);
            t.dst[3] = hi32(d);
        }
        const uint32_t npfx = t.ver == 6 ? s.n_pfx6 : s.n_pfx4;
        if (npfx && hi32(q) % 1000u < s.lpm_permille) {
            uint32_t w[4], port;
            int len;
            prefix_of(s, (int)t.ver, hi32(finalise(q)) % npfx, w, len, port);
            for (int k = 0; k < (t.ver == 6 ? 4 : 1); ++k) {
                const uint32_t m = top_bits(len - 32 * k);
                t.dst[k] = (w[k] & m) | (t.dst[k] & ~m);
            }
        }
    }
    int at = 12;
    if (t.vlan) {
        o.u16(12, 0x8100);
        o.u16(14, t.vlan | 0x6000u);  // PCP 3: a parser must mask the VID
        at = 16;
    }
    const uint32_t quirk = (uint32_t)(p3 >> 56) & 0xFFu;
    if (t.ver == 4) {
        const uint32_t ihl = quirk == 0 ? 6 : 5;  // one packet in 256 carries four NOP options
        const int ip = at + 2;
        o.u16(at, 0x0800);
        o.u8(ip, 0x40u | ihl);
        o.u16(ip + 2, 46);
        o.u16(ip + 4, (uint32_t)(p3 >> 24));
        o.u16(ip + 6, quirk == 0x11u ? 185u : 0x4000u);  // one in 256 is a non-first fragment, the rest set DF
        o.u8(ip + 8, 64);
        o.u8(ip + 9, t.proto);
        o.u32(ip + 12, t.src[0]);
        o.u32(ip + 16, t.dst[0]);
        if (ihl == 6) o.u32(ip + 20, 0x01010101u);
        o.u16(ip + 4 * (int)ihl, t.sport);
        o.u16(ip + 4 * (int)ihl + 2, t.dport);
    } else {
        const int ip = at + 2;
        o.u16(at, 0x86DD);
        o.u8(ip, 0x60);
        o.u16(ip + 4, 8);
        o.u8(ip + 6, t.proto);
        o.u8(ip + 7, 64);
        for (int k = 0; k < 4; ++k) {
            o.u32(ip + 8 + 4 * k, t.src[k]);
            o.u32(ip + 24 + 4 * k, t.dst[k]);
        }
        o.u16(ip + 40, t.sport);
        o.u16(ip + 42, t.dport);
    }
}

void put_be32(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)(v >> 24);
    p[1] = (uint8_t)(v >> 16);
    p[2] = (uint8_t)(v >> 8);
    p[3] = (uint8_t)v;
}

}  // namespace

extern "C" {

void orc_gen_packets(const sfc_gen_spec* spec, uint64_t first, size_t n, uint8_t* hdr, uint16_t* in_port) {
    for (size_t i = 0; i < n; ++i) packet_of(*spec, first + i, hdr + 64 * i, in_port[i]);
}

void orc_gen_flows(const sfc_gen_spec* spec, uint32_t first, size_t n, sfc_flow_key* keys, sfc_flow_action* acts) {
    for (size_t i = 0; i < n; ++i) {
        const Tuple t = flow_of(*spec, first + (uint32_t)i);
        memset(&keys[i], 0, sizeof(sfc_flow_key));
        for (int k = 0; k < (t.ver == 6 ? 4 : 1); ++k) {
            put_be32(keys[i].src + 4 * k, t.src[k]);
            put_be32(keys[i].dst + 4 * k, t.dst[k]);
        }
        keys[i].sport = (uint16_t)t.sport;
        keys[i].dport = (uint16_t)t.dport;
        keys[i].vlan = (uint16_t)t.vlan;
        keys[i].proto = (uint8_t)t.proto;
        keys[i].ip_ver = (uint8_t)t.ver;
        acts[i].action = (uint8_t)t.action;
        acts[i].reserved = 0;
        acts[i].out_port = (uint16_t)t.out_port;
    }
}

void orc_gen_mac(uint32_t i, uint8_t* mac) { mac_of(i, mac); }

void orc_gen_prefix(const sfc_gen_spec* spec, int af, uint32_t j, uint8_t* prefix, int* len, uint16_t* out_port) {
    uint32_t w[4], port;
    int l;
    prefix_of(*spec, af, j, w, l, port);
    for (int k = 0; k < (af == 6 ? 4 : 1); ++k) put_be32(prefix + 4 * k, w[k]);
    *len = l;
    *out_port = (uint16_t)port;
}

// ACL rule j: an IPv4 (80 %) or IPv6 rule whose destination is a prefix of the LPM set cut to a coarse length,
// a mostly-open source, optional protocol / destination-port / in_port constraints, one rule in ten denying;
// priority grows with the number of bits the rule fixes (ties on purpose).
void orc_gen_acl_rule(const sfc_gen_spec* spec, uint32_t j, sfc_acl_rule* r) {
    const sfc_gen_spec& s = *spec;
    memset(r, 0, sizeof(*r));
    const uint64_t a0 = R(s.seed_acl, 4ull * j), a1 = R(s.seed_acl, 4ull * j + 1), a2 = R(s.seed_acl, 4ull * j + 2),
                   a3 = R(s.seed_acl, 4ull * j + 3);
    const bool six = s.n_pfx6 && lo32(a0) % 10u < 2u;
    r->ip_ver = six ? 6 : 4;
    uint32_t w[4] = {hi32(a1), lo32(a1), hi32(a2), lo32(a2)}, port;
    int plen = six ? 64 : 32;
    const uint32_t npfx = six ? s.n_pfx6 : s.n_pfx4;
    if (npfx) prefix_of(s, six ? 6 : 4, hi32(a1) % npfx, w, plen, port);
    static const int cut4[4] = {8, 16, 24, 32}, cut6[4] = {32, 40, 48, 64};
    int dlen = (six ? cut6 : cut4)[(a0 >> 8) & 3];
    if (dlen > plen) dlen = plen;
    if (((a0 >> 12) & 7) == 0) dlen = 0;  // one in eight: any destination
    for (int k = 0; k < (six ? 4 : 1); ++k) {
        const uint32_t m = top_bits(dlen - 32 * k);
        put_be32(r->dst + 4 * k, w[k] & m);
        put_be32(r->dst_mask + 4 * k, m);
    }
    const uint32_t sl = (uint32_t)(a0 >> 16) & 15u;
    const int slen = sl < 11 ? 0 : sl < 14 ? 4 : 8;
    put_be32(r->src, hi32(a2) & top_bits(slen));
    put_be32(r->src_mask, top_bits(slen));
    if ((a0 >> 20) & 1) {
        r->proto = ((a0 >> 21) & 1) ? 17 : 6;
        r->proto_mask = 0xFF;
    }
    const uint32_t dp = (uint32_t)(a0 >> 24) & 15u;
    if (dp < 5) {  // an exact well-known port
        r->dport = (uint16_t)((a3 >> 8) & 0x3FF);
        r->dport_mask = 0xFFFF;
    } else if (dp < 8) {  // a block of 4096 ports
        r->dport = (uint16_t)((a3 >> 8) & 0xF000);
        r->dport_mask = 0xF000;
    }
    if (((a0 >> 28) & 15u) == 0) {
        r->in_port = (uint16_t)(s.port_base + ((uint32_t)(a3 >> 24) & 0xFFFFu) % (s.n_ports ? s.n_ports : 1u));
        r->in_port_mask = 0xFFFF;
    }
    r->verdict = lo32(a3) % 10u == 0 ? SFC_ACL_DENY : SFC_ACL_PERMIT;
    const uint32_t fixed_bits = (uint32_t)dlen + (r->dport_mask == 0xFFFF ? 16u : r->dport_mask ? 4u : 0u) + (uint32_t)slen;
    r->priority = 1u + fixed_bits * 4u + hi32(a3) % 4u;
}

}  // extern "C"
