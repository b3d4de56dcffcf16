// sfc_oracle.cpp — CPU oracle for the SFC flow-classification path.  TEST INFRASTRUCTURE ONLY:
// nothing in the product library (dpu-operator_b200/) includes, links or calls this file; only
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs load it.
//
// PARITY UNPINNED at per-packet granularity.  The reference (Ximinhan/dpu-operator) is a Go
// control plane that never classifies a packet itself — OVS or DPU hardware does — and not one
// of its tests asserts a rule string, a table entry or a forwarding decision.
// There is therefore no golden per-packet vector to pin this file against.  What it restates, with
// the reference lines it follows, is the *meaning* of the rules the reference emits:
//   - L2 steering rule shapes and their OpenFlow priorities: R1 "priority=10,in_port=S ->
//     output:D", R2 "in_port=S,dl_dst=M -> output:D" (default priority 32768), R3 hairpin
//     "priority=100,in_port=S,dl_dst=M -> in_port"
//     (internal/daemon/vendor-specific-plugins/marvell/ovs-dp/ovsdp.go:113-142); OpenFlow picks the
//     highest-priority matching flow, so MAC-qualified rules beat the in_port wildcard.
//   - ternary match + priority: P4Runtime semantics as p4rt-ctl encodes them
//     (cmd/intelvsp/p4rt-ctl:380-404 match kinds, :455-534 priority required for ternary);
//     higher number wins.
//   - LPM: longest prefix wins (ipv4_table / ipv6_table,
//     cmd/intelvsp/fxp-net_linux-networking/fxp-net_linux-networking.p4info.txt:807,868).
//   - miss -> punt to control plane (default actions fwd_to_cp / fwd_to_cp_tx of l2_fwd_rx_table /
//     l2_fwd_tx_table in fxp-net_linux-networking.context.json).
//   - wire byte order of the header fields: big-endian, as in dpu-cni/pkgs/sriovutils/packet.go:41-66.
// The pinned stage order is SURVEY.md §8a "Proposed pinned semantics".  The rule *sets* are pinned
// separately against the reference's generators in tests/test_replay.py.
//
// Deliberately plain: byte-at-a-time parser, std::unordered_map over the 40 key bytes for exact matches, a linear
// priority scan for the ACL, a one-bit-per-level binary trie for LPM.  It shares no code with the
// CUDA library (only the public structs of include/b200sfc.h).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../include/b200sfc.h"

namespace {

struct Parsed {
    uint8_t dmac[6];
    uint16_t vlan;
    uint8_t ip_ver;  // 0 = not IP
    uint8_t proto;
    uint8_t src[16], dst[16];
    uint16_t sport, dport;
};

inline uint16_t rd16(const uint8_t* p) { return (uint16_t)((p[0] << 8) | p[1]); }

// stage 0 — byte order as on the wire (dpu-cni/pkgs/sriovutils/packet.go:41-66)
void parse(const uint8_t* b, Parsed& k) {
    memset(&k, 0, sizeof(k));
    memcpy(k.dmac, b, 6);
    size_t off = 12;
    uint16_t et = rd16(b + off);
    off += 2;
    if (et == 0x8100) {  // one 802.1Q tag
        k.vlan = rd16(b + off) & 0x0FFF;
        et = rd16(b + off + 2);
        off += 4;
    }
    if (et == 0x0800) {
        const uint8_t* ip = b + off;
        unsigned ver = ip[0] >> 4, ihl = ip[0] & 0x0F;
        if (ver != 4 || ihl < 5) return;
        k.ip_ver = 4;
        k.proto = ip[9];
        memcpy(k.src, ip + 12, 4);
        memcpy(k.dst, ip + 16, 4);
        unsigned frag_off = rd16(ip + 6) & 0x1FFF;
        size_t l4 = off + 4 * ihl;
        bool l4proto = k.proto == 6 || k.proto == 17 || k.proto == 132;
        if (l4proto && frag_off == 0 && l4 + 4 <= SFC_HDR_BYTES) {
            k.sport = rd16(b + l4);
            k.dport = rd16(b + l4 + 2);
        }
    } else if (et == 0x86DD) {
        const uint8_t* ip = b + off;
        if ((ip[0] >> 4) != 6) return;
        k.ip_ver = 6;
        k.proto = ip[6];
        memcpy(k.src, ip + 8, 16);
        memcpy(k.dst, ip + 24, 16);
        size_t l4 = off + 40;
        bool l4proto = k.proto == 6 || k.proto == 17 || k.proto == 132;
        if (l4proto && l4 + 4 <= SFC_HDR_BYTES) {
            k.sport = rd16(b + l4);
            k.dport = rd16(b + l4 + 2);
        }
    }
}

struct L2Rule {
    uint16_t in_port, out_port;
    bool has_mac;
    uint8_t mac[6];
    uint32_t priority, id;
};
struct FlowVal {
    uint8_t action;
    uint16_t out_port;
    uint32_t id;
};
struct AclRule {
    sfc_acl_rule r;
    uint32_t id;
};
struct TrieNode {
    int32_t child[2] = {-1, -1};
    bool has = false;
    uint32_t id = 0;
    uint16_t port = 0;
};
struct BitTrie {
    std::vector<TrieNode> n = std::vector<TrieNode>(1);
    void add(const uint8_t* p, int len, uint16_t port, uint32_t id) {
        int cur = 0;
        for (int i = 0; i < len; ++i) {
            int bit = (p[i >> 3] >> (7 - (i & 7))) & 1;
            if (n[cur].child[bit] < 0) {
                n[cur].child[bit] = (int32_t)n.size();
                n.emplace_back();
            }
            cur = n[cur].child[bit];
        }
        n[cur].has = true;
        n[cur].id = id;
        n[cur].port = port;
    }
    bool longest(const uint8_t* a, int bits, uint16_t& port, uint32_t& id) const {
        int cur = 0;
        bool found = false;
        for (int i = 0;; ++i) {
            if (n[cur].has) {
                found = true;
                port = n[cur].port;
                id = n[cur].id;
            }
            if (i == bits) break;
            int bit = (a[i >> 3] >> (7 - (i & 7))) & 1;
            if (n[cur].child[bit] < 0) break;
            cur = n[cur].child[bit];
        }
        return found;
    }
};

// exact-match key: the 40 bytes (ip_ver, proto, vlan, sport, dport, src, dst) as they are, hashed with FNV-1a
struct FlowKey {
    uint8_t b[40];
    bool operator==(const FlowKey& o) const { return !memcmp(b, o.b, 40); }
};
struct FlowKeyHash {
    size_t operator()(const FlowKey& k) const {
        uint64_t h = 0xcbf29ce484222325ull;
        for (int i = 0; i < 40; ++i) h = (h ^ k.b[i]) * 0x100000001b3ull;
        return (size_t)(h ^ (h >> 29));
    }
};
FlowKey flow_key_bytes(uint8_t ip_ver, uint16_t vlan, uint8_t proto, const uint8_t* src, const uint8_t* dst,
                       uint16_t sport, uint16_t dport) {
    FlowKey s;
    s.b[0] = ip_ver;
    s.b[1] = proto;
    s.b[2] = (uint8_t)(vlan >> 8);
    s.b[3] = (uint8_t)vlan;
    s.b[4] = (uint8_t)(sport >> 8);
    s.b[5] = (uint8_t)sport;
    s.b[6] = (uint8_t)(dport >> 8);
    s.b[7] = (uint8_t)dport;
    memcpy(s.b + 8, src, 16);
    memcpy(s.b + 24, dst, 16);
    return s;
}

}  // namespace

struct oracle {
    std::vector<L2Rule> l2;
    std::vector<uint16_t> nf_in, nf_out;
    std::unordered_map<FlowKey, FlowVal, FlowKeyHash> flows;
    BitTrie lpm4, lpm6;
    bool has_lpm4 = false, has_lpm6 = false;
    std::vector<AclRule> acl;

    uint8_t nf_of(uint16_t port) const {
        for (size_t i = 0; i < nf_in.size(); ++i)
            if (nf_in[i] == port || nf_out[i] == port) return (uint8_t)i;
        return SFC_NF_NONE;
    }

    uint64_t classify_one(const uint8_t* hdr, uint16_t in_port, uint64_t* counters) const {
        Parsed k;
        parse(hdr, k);
        uint64_t res = SFC_RESULT_MISS;
        bool done = false;
        // stage 1: ACL, highest priority wins, ties -> lowest id
        if (k.ip_ver) {
            const AclRule* best = nullptr;
            for (auto& a : acl) {
                const sfc_acl_rule& r = a.r;
                if (r.ip_ver && r.ip_ver != k.ip_ver) continue;
                if ((in_port & r.in_port_mask) != (r.in_port & r.in_port_mask)) continue;
                if ((k.vlan & r.vlan_mask) != (r.vlan & r.vlan_mask)) continue;
                if ((k.proto & r.proto_mask) != (r.proto & r.proto_mask)) continue;
                if ((k.sport & r.sport_mask) != (r.sport & r.sport_mask)) continue;
                if ((k.dport & r.dport_mask) != (r.dport & r.dport_mask)) continue;
                bool ok = true;
                for (int i = 0; i < 16 && ok; ++i)
                    ok = ((k.src[i] & r.src_mask[i]) == (r.src[i] & r.src_mask[i])) &&
                         ((k.dst[i] & r.dst_mask[i]) == (r.dst[i] & r.dst_mask[i]));
                if (!ok) continue;
                if (!best || r.priority > best->r.priority || (r.priority == best->r.priority && a.id < best->id))
                    best = &a;
            }
            if (best) {
                if (counters) counters[SFC_CNT_RULES0 + best->id]++;
                if (best->r.verdict == SFC_ACL_DENY) {
                    res = SFC_RESULT(SFC_ACT_DROP, SFC_NF_NONE, 0, SFC_STAGE_ACL, best->id);
                    done = true;
                }
            }
        }
        // stage 2: exact 5-tuple
        if (!done && k.ip_ver) {
            auto it = flows.find(flow_key_bytes(k.ip_ver, k.vlan, k.proto, k.src, k.dst, k.sport, k.dport));
            if (it != flows.end()) {
                const FlowVal& v = it->second;
                uint8_t nf = v.action == SFC_ACT_FORWARD ? nf_of(v.out_port) : (uint8_t)SFC_NF_NONE;
                res = SFC_RESULT(v.action, nf, v.out_port, SFC_STAGE_FLOW, v.id);
                done = true;
            }
        }
        // stage 3: L2 steering — OpenFlow: highest priority among matching flows
        // (ovs-dp/ovsdp.go:120 prio 100, :127 default 32768, :135 prio 10)
        if (!done) {
            const L2Rule* best = nullptr;
            for (auto& r : l2) {
                if (r.in_port != in_port) continue;
                if (r.has_mac && memcmp(r.mac, k.dmac, 6)) continue;
                if (!best || r.priority > best->priority || (r.priority == best->priority && r.id < best->id))
                    best = &r;
            }
            if (best) {
                res = SFC_RESULT(SFC_ACT_FORWARD, nf_of(best->out_port), best->out_port,
                                 best->has_mac ? SFC_STAGE_L2_EXACT : SFC_STAGE_L2_WILD, best->id);
                if (counters) counters[SFC_CNT_L2_0 + best->id % SFC_CNT_L2_N]++;
                done = true;
            }
        }
        // stage 4: LPM on the destination
        if (!done && k.ip_ver) {
            uint16_t port = 0;
            uint32_t id = 0;
            if (k.ip_ver == 4 && has_lpm4 && lpm4.longest(k.dst, 32, port, id))
                res = SFC_RESULT(SFC_ACT_FORWARD, nf_of(port), port, SFC_STAGE_LPM4, id);
            else if (k.ip_ver == 6 && has_lpm6 && lpm6.longest(k.dst, 128, port, id))
                res = SFC_RESULT(SFC_ACT_FORWARD, nf_of(port), port, SFC_STAGE_LPM6, id);
        }
        if (counters) {
            counters[SFC_CNT_ACTION0 + SFC_RESULT_ACTION(res)]++;
            counters[SFC_CNT_STAGE0 + SFC_RESULT_STAGE(res)]++;
        }
        return res;
    }
};

extern "C" {

oracle* orc_create(void) { return new oracle(); }
void orc_destroy(oracle* o) { delete o; }
void orc_clear(oracle* o) { *o = oracle(); }

// priority is the OpenFlow priority the reference would have written (10 / 100 / 32768)
void orc_l2_add(oracle* o, uint16_t in_port, uint16_t out_port, const uint8_t* mac, uint32_t priority, uint32_t id) {
    L2Rule r{};
    r.in_port = in_port;
    r.out_port = out_port;
    r.has_mac = mac != nullptr;
    if (mac) memcpy(r.mac, mac, 6);
    r.priority = priority;
    r.id = id;
    // add-flow with an identical match and priority replaces the actions (ovs-ofctl semantics)
    for (auto& e : o->l2)
        if (e.in_port == in_port && e.has_mac == r.has_mac && e.priority == priority &&
            (!mac || !memcmp(e.mac, mac, 6))) {
            e.out_port = out_port;
            return;
        }
    o->l2.push_back(r);
}
// del-flows in_port=S[,dl_dst=M] is a loose match (ovs-dp/ovsdp.go:143-162)
void orc_l2_del(oracle* o, uint16_t in_port, const uint8_t* mac) {
    auto& v = o->l2;
    v.erase(std::remove_if(v.begin(), v.end(),
                           [&](const L2Rule& e) {
                               return e.in_port == in_port && (!mac || (e.has_mac && !memcmp(e.mac, mac, 6)));
                           }),
            v.end());
}
void orc_chain_set(oracle* o, const uint16_t* in, const uint16_t* out, size_t n) {
    o->nf_in.assign(in, in + n);
    o->nf_out.assign(out, out + n);
}
void orc_flow_add(oracle* o, const sfc_flow_key* k, const sfc_flow_action* a, size_t n, uint32_t first_id) {
    o->flows.reserve(o->flows.size() + n);
    for (size_t i = 0; i < n; ++i)
        o->flows[flow_key_bytes(k[i].ip_ver, k[i].vlan, k[i].proto, k[i].src, k[i].dst, k[i].sport, k[i].dport)] =
            FlowVal{a[i].action, a[i].out_port, first_id + (uint32_t)i};
}
void orc_flow_del(oracle* o, const sfc_flow_key* k) {
    o->flows.erase(flow_key_bytes(k->ip_ver, k->vlan, k->proto, k->src, k->dst, k->sport, k->dport));
}
void orc_lpm_add(oracle* o, int af, const uint8_t* prefix, int len, uint16_t port, uint32_t id) {
    (af == 4 ? o->lpm4 : o->lpm6).add(prefix, len, port, id);
    (af == 4 ? o->has_lpm4 : o->has_lpm6) = true;
}
void orc_acl_add(oracle* o, const sfc_acl_rule* r, uint32_t id) { o->acl.push_back(AclRule{*r, id}); }

// classify n packets on `threads` host threads (contiguous split); counters may be NULL, else they are
// ADDED to (each thread counts into its own array, the arrays are summed at the end)
void orc_classify(const oracle* o, const uint8_t* hdr, const uint16_t* in_port, uint64_t* result, size_t n,
                  int threads, uint64_t* counters) {
    if (threads <= 1) {
        for (size_t i = 0; i < n; ++i) result[i] = o->classify_one(hdr + i * SFC_HDR_BYTES, in_port[i], counters);
        return;
    }
    std::vector<std::thread> pool;
    std::vector<std::vector<uint64_t>> cnt(counters ? threads : 0, std::vector<uint64_t>(SFC_CNT_TOTAL, 0));
    size_t per = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        size_t lo = std::min(n, per * t), hi = std::min(n, lo + per);
        uint64_t* mine = counters ? cnt[t].data() : nullptr;
        pool.emplace_back([=] {
            for (size_t i = lo; i < hi; ++i) result[i] = o->classify_one(hdr + i * SFC_HDR_BYTES, in_port[i], mine);
        });
    }
    for (auto& th : pool) th.join();
    for (auto& c : cnt)
        for (size_t i = 0; i < SFC_CNT_TOTAL; ++i) counters[i] += c[i];
}

// expose the parser for unit tests: out = {vlan, ip_ver, proto, sport, dport}, addresses, dmac
void orc_parse(const uint8_t* hdr, uint32_t* fields5, uint8_t* src16, uint8_t* dst16, uint8_t* dmac6) {
    Parsed k;
    parse(hdr, k);
    fields5[0] = k.vlan;
    fields5[1] = k.ip_ver;
    fields5[2] = k.proto;
    fields5[3] = k.sport;
    fields5[4] = k.dport;
    memcpy(src16, k.src, 16);
    memcpy(dst16, k.dst, 16);
    memcpy(dmac6, k.dmac, 6);
}

}  // extern "C"
