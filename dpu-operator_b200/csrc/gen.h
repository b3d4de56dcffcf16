// gen.h — deterministic synthetic workload: flows, prefixes, ACL rules and
// packet headers are pure functions of (seed, index) through counter-based splitmix64, so the
// host and a CUDA kernel produce identical bytes from this single source.  Integer-only.
#pragma once
#include "../../include/b200sfc.h"
#include "sfc_common.h"

namespace sfc {

struct GenFlow {
    uint32_t sip[4], dip[4];
    uint32_t sport, dport, vlan, proto, ip_ver;
    uint32_t action, out_port;
};

SFC_HD void gen_mac(uint32_t i, uint8_t* mac) {
    // the reference's VF MAC pattern 00:<vsi>:00:00:03:14 (vendored ipuplugin/bridgeport.go:99-100)
    mac[0] = 0x00;
    mac[1] = (uint8_t)i;
    mac[2] = (uint8_t)(i >> 8);
    mac[3] = (uint8_t)(i >> 16);
    mac[4] = 0x03;
    mac[5] = 0x14;
}

SFC_HD uint32_t gen_vlan(uint64_t r, uint32_t vlan_permille) {
    // bits 1..10 decide, bits 12..23 pick the VID
    uint32_t t = (uint32_t)(r >> 1) & 1023u;
    bool tagged = ((t * 1000u) >> 10) < vlan_permille;
    return tagged ? 1u + ((uint32_t)(r >> 12) & 0xFFFu) % 4094u : 0u;
}

SFC_HD void gen_flow(const sfc_gen_spec& s, uint32_t i, GenFlow& f) {
    uint64_t r0 = draw(s.seed_rules, 4ull * i + 0), r1 = draw(s.seed_rules, 4ull * i + 1);
    uint64_t r2 = draw(s.seed_rules, 4ull * i + 2), r3 = draw(s.seed_rules, 4ull * i + 3);
    bool v6 = ((uint32_t)(r0 >> 32) % 1000u) < s.v6_permille;
    f.ip_ver = v6 ? 6u : 4u;
    f.proto = (r0 & 1) ? 17u : 6u;
    f.vlan = gen_vlan(r0, s.vlan_permille);
    f.sport = (uint32_t)r1 & 0xFFFFu;
    f.dport = (uint32_t)(r1 >> 16) & 0xFFFFu;
    bool drop = ((uint32_t)(r1 >> 32) & 0xFFu) < 13u;
    f.action = drop ? SFC_ACT_DROP : SFC_ACT_FORWARD;
    uint32_t nout = s.n_out_ports ? s.n_out_ports : 1u;
    f.out_port = drop ? 0u : ((uint32_t)(r1 >> 40) & 0xFFFFu) % nout;
    for (int k = 0; k < 4; ++k) f.sip[k] = f.dip[k] = 0;
    if (!v6) {
        f.sip[0] = (uint32_t)(r2 >> 32);
        f.dip[0] = (uint32_t)r2;
    } else {
        uint64_t d0 = mix64(r2 ^ 0x6A09E667F3BCC909ull), d1 = mix64(r3 ^ 0xBB67AE8584CAA73Bull);
        f.sip[0] = (uint32_t)(r2 >> 32);
        f.sip[1] = (uint32_t)r2;
        f.sip[2] = (uint32_t)(r3 >> 32);
        f.sip[3] = (uint32_t)r3;
        f.dip[0] = (uint32_t)(d0 >> 32);
        f.dip[1] = (uint32_t)d0;
        f.dip[2] = (uint32_t)(d1 >> 32);
        f.dip[3] = (uint32_t)d1;
    }
}

// BGP-like IPv4 prefix-length histogram (per mille), peak /24, range /8../32
SFC_HD int gen_len4(uint32_t u) {
    const uint16_t cum[25] = {1,   2,   3,   4,   6,   9,   14,  19,  59,  74,  94,  129, 174,
                              219, 309, 379, 939, 944, 949, 954, 962, 972, 984, 985, 1000};
    for (int k = 0; k < 25; ++k)
        if (u < cum[k]) return 8 + k;
    return 32;
}
// IPv6: /32../64, peak /48
SFC_HD int gen_len6(uint32_t u) {
    const uint16_t cum[9] = {100, 130, 190, 250, 750, 770, 870, 900, 1000};
    const uint8_t len[9] = {32, 36, 40, 44, 48, 52, 56, 60, 64};
    for (int k = 0; k < 9; ++k)
        if (u < cum[k]) return len[k];
    return 64;
}
SFC_HD uint32_t mask32(int len) {  // top `len` bits of a word, len in 0..32
    return len <= 0 ? 0u : (len >= 32 ? 0xFFFFFFFFu : ~(0xFFFFFFFFu >> len));
}
// prefix j: words w[4] (masked), length, next-hop port
SFC_HD void gen_prefix(const sfc_gen_spec& s, int af, uint32_t j, uint32_t* w, int& len,
                       uint32_t& port) {
    uint64_t seed = s.seed_lpm ^ (af == 6 ? 0x6666666666666666ull : 0ull);
    uint64_t r = draw(seed, 2ull * j), q = draw(seed, 2ull * j + 1);
    uint32_t nout = s.n_out_ports ? s.n_out_ports : 1u;
    port = ((uint32_t)q & 0xFFFFu) % nout;
    w[0] = w[1] = w[2] = w[3] = 0;
    if (af == 4) {
        len = gen_len4((uint32_t)r % 1000u);
        w[0] = (uint32_t)(r >> 32) & mask32(len);
    } else {
        len = gen_len6((uint32_t)r % 1000u);
        uint32_t a = 0x20000000u | ((uint32_t)(r >> 32) & 0x1FFFFFFFu);  // 2000::/3
        uint32_t b = (uint32_t)(q >> 32);
        w[0] = a & mask32(len);
        w[1] = b & mask32(len - 32);
    }
}

SFC_HD uint32_t pick_flow(const sfc_gen_spec& s, uint64_t p1) {
    if (!s.zipf) return (uint32_t)(p1 % s.n_flows);
    // octave k covers ranks [2^k - 1, 2^(k+1) - 1) and every octave has the same mass
    uint32_t L = 31u - (uint32_t)
#if defined(__CUDA_ARCH__)
                           __clz((int)s.n_flows);
#else
                           __builtin_clz(s.n_flows);
#endif
    uint32_t k = (uint32_t)(p1 >> 58) % (L + 1u);
    uint64_t base = (1ull << k) - 1ull;
    uint64_t f = base + ((p1 & ((1ull << 58) - 1ull)) & ((1ull << k) - 1ull));
    return (uint32_t)(f % s.n_flows);
}

SFC_HD void put16(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)(v >> 8);
    p[1] = (uint8_t)v;
}
SFC_HD void put32(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)(v >> 24);
    p[1] = (uint8_t)(v >> 16);
    p[2] = (uint8_t)(v >> 8);
    p[3] = (uint8_t)v;
}

// one 64-byte header record + in_port
SFC_HD void gen_packet(const sfc_gen_spec& s, uint64_t i, uint8_t* h, uint16_t& in_port) {
    uint64_t p0 = draw(s.seed_pkts, 4 * i + 0), p1 = draw(s.seed_pkts, 4 * i + 1);
    uint64_t p2 = draw(s.seed_pkts, 4 * i + 2), p3 = draw(s.seed_pkts, 4 * i + 3);
    for (int k = 0; k < 64; ++k) h[k] = 0;
    uint32_t nports = s.n_ports ? s.n_ports : 1u, nmacs = s.n_macs ? s.n_macs : 1u;
    in_port = (uint16_t)(s.port_base + ((uint32_t)p3 & 0xFFFFu) % nports);
    gen_mac(((uint32_t)(p3 >> 16) & 0xFFFFFFu) % nmacs, h);
    h[6] = 0x02;
    h[9] = (uint8_t)(p3 >> 40);
    h[10] = (uint8_t)(p3 >> 48);
    h[11] = (uint8_t)(p2 >> 8);

    uint32_t c = (uint32_t)(p0 >> 32) % 1000u;
    if (c < s.nonip_permille) {
        uint32_t vlan = gen_vlan(p1, s.vlan_permille);
        int off = 12;
        if (vlan) {
            put16(h + 12, 0x8100);
            put16(h + 14, vlan | 0x2000u);  // PCP 1: the parser must mask the VID
            off = 16;
        }
        put16(h + off, 0x0806);
        put32(h + off + 2, 0x00010800u);
        put32(h + off + 6, 0x06040001u);
        put32(h + off + 10, (uint32_t)(p1 >> 32));
        put32(h + off + 14, (uint32_t)p2);
        put32(h + off + 18, (uint32_t)(p2 >> 32));
        return;
    }
    GenFlow f;
    bool hit = s.n_flows > 0 && ((uint32_t)p0 % 1000u) < s.hit_permille;
    if (hit) {
        gen_flow(s, pick_flow(s, p1), f);
    } else {
        uint64_t q = mix64(p1);
        bool v6 = ((uint32_t)(p1 >> 32) % 1000u) < s.v6_permille;
        f.ip_ver = v6 ? 6u : 4u;
        const uint8_t protos[4] = {6, 17, 1, 132};
        f.proto = protos[p1 & 3];
        f.vlan = gen_vlan(p1 >> 1, s.vlan_permille);
        f.sport = (uint32_t)q & 0xFFFFu;
        f.dport = (uint32_t)(q >> 16) & 0xFFFFu;
        for (int k = 0; k < 4; ++k) f.sip[k] = f.dip[k] = 0;
        uint64_t a = mix64(p2), b = mix64(p2 ^ 0x3C6EF372FE94F82Bull), d = mix64(a ^ b);
        f.sip[0] = (uint32_t)(p2 >> 32);
        f.dip[0] = (uint32_t)p2;
        if (v6) {
            f.sip[1] = (uint32_t)a;
            f.sip[2] = (uint32_t)(a >> 32);
            f.sip[3] = (uint32_t)b;
            f.dip[1] = (uint32_t)(b >> 32);
            f.dip[2] = (uint32_t)d;
            f.dip[3] = (uint32_t)(d >> 32);
        }
        uint32_t npfx = v6 ? s.n_pfx6 : s.n_pfx4;
        if (npfx && ((uint32_t)(q >> 32) % 1000u) < s.lpm_permille) {
            uint32_t w[4], port;
            int len;
            gen_prefix(s, v6 ? 6 : 4, (uint32_t)(mix64(q) >> 32) % npfx, w, len, port);
            for (int k = 0; k < 4; ++k) {
                uint32_t m = mask32(len - 32 * k);
                f.dip[k] = (w[k] & m) | (f.dip[k] & ~m);
                if (!v6) break;
            }
        }
    }
    int off = 12;
    if (f.vlan) {
        put16(h + 12, 0x8100);
        put16(h + 14, f.vlan | 0x6000u);  // PCP 3
        off = 16;
    }
    if (f.ip_ver == 4) {
        put16(h + off, 0x0800);
        uint8_t* ip = h + off + 2;
        uint32_t ihl = ((uint32_t)(p3 >> 56) & 0xFFu) == 0 ? 6u : 5u;  // 1 in 256 carries IP options
        bool frag = ((uint32_t)(p3 >> 56) & 0xFFu) == 0x11u;
        ip[0] = (uint8_t)(0x40u | ihl);
        put16(ip + 2, 46);
        put16(ip + 4, (uint32_t)(p3 >> 24));
        put16(ip + 6, frag ? 185u : 0x4000u);  // DF set, offset 0 | fragment at offset 185
        ip[8] = 64;
        ip[9] = (uint8_t)f.proto;
        put32(ip + 12, f.sip[0]);
        put32(ip + 16, f.dip[0]);
        if (ihl == 6) put32(ip + 20, 0x01010101u);  // four NOP options
        put16(ip + 4 * ihl, f.sport);
        put16(ip + 4 * ihl + 2, f.dport);
    } else {
        put16(h + off, 0x86DD);
        uint8_t* ip = h + off + 2;
        ip[0] = 0x60;
        put16(ip + 4, 8);
        ip[6] = (uint8_t)f.proto;
        ip[7] = 64;
        for (int k = 0; k < 4; ++k) {
            put32(ip + 8 + 4 * k, f.sip[k]);
            put32(ip + 24 + 4 * k, f.dip[k]);
        }
        put16(ip + 40, f.sport);
        put16(ip + 42, f.dport);
    }
}

}  // namespace sfc
