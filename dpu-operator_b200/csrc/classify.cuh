// classify.cuh — the per-packet match-action pipeline as one persistent sm_100a kernel.
//
// Data movement: header records stream HBM -> shared memory through TMA (cp.async.bulk.tensor,
// 64-byte swizzle so that "thread p reads packet p" is bank-conflict free), a ring of STAGES
// tiles per CTA guarded by full/empty mbarriers, one producer lane and TILE consumer threads.
// Each consumer owns one packet of a tile: 4 x LDS.128 -> registers, slot released at once,
// then parse -> ACL -> flow probe -> L2 steering -> LPM, and one 8-byte result store.
// Tables are read through the read-only global path (one 32-byte sector per flow-table probe).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200sfc.h"
#include "sfc_common.h"

namespace sfc {

constexpr uint32_t kProbeEager = 1u, kProbeL1 = 2u;

struct DevTables {
    const uint32_t* flow4;
    const uint32_t* flow6;
    uint32_t flow4_nbuckets, flow4_nhash, flow4_count;
    uint32_t flow4_mode;  // kProbeEager: issue both probes at once; kProbeL1: let probes allocate in L1
    uint32_t flow4_slots, flow6_slots;  // slots per bucket (1 or 2)
    uint32_t flow6_nbuckets, flow6_nhash, flow6_count;
    uint32_t S4[kFlow4SeedWords];
    uint32_t S6[kFlow6SeedWords];
    // format 2 (indexed): one fingerprint index + dense record units for both families
    const uint32_t* fidx;
    const uint32_t* frec;
    uint32_t flow_format, fidx_nbuckets, fidx_bits;
    uint32_t S2[kFlow6SeedWords];
    const L2Entry* l2x;
    uint32_t l2x_mask, l2x_count;
    const L2Wild* l2w;
    uint32_t l2w_n, l2w_count;
    const uint32_t* lpm4_root;
    const uint32_t* lpm4_nodes;
    const LpmRes* lpm4_nh;
    const LpmTail* lpm4_tails;
    uint32_t lpm4_count;
    const uint32_t* lpm6_root;
    const uint32_t* lpm6_nodes;
    const LpmRes* lpm6_nh;
    const LpmTail* lpm6_tails;
    uint32_t lpm6_count;
    const uint4* acl_nodes;               // AclEntry arrays, 2^bits per node
    const uint16_t* acl_leaves;           // overflow candidate lists
    const uint32_t* acl_rules;            // kAclRuleWords per rank: groups G0, G1, G2
    const uint32_t* acl_res;              // per rank: rule id | deny << 31
    uint32_t acl_nrules, acl_ntrees;
    uint4 acl_roots[kAclMaxTrees];
    unsigned long long* counters;
};

// One launch over several independent packet batches (sfc_classify_batches).  The list travels as a
// kernel parameter: nothing to allocate or copy, and a captured launch carries its own copy.
constexpr int kMaxBatches = 512;
struct BatchDesc {
    uint32_t row0;   // first record of the batch, in 64-byte rows from BatchList::base
    uint32_t n;      // packets in the batch
    uint32_t tile0;  // index of the batch's first tile within the launch
    uint32_t pad;
    const uint16_t* in_port;
    uint64_t* result;
};
struct BatchList {
    const uint8_t* base;  // lowest header address of the launch; the tensor map starts here
    uint32_t count, pad;
    BatchDesc b[kMaxBatches];
};
struct NoBatches {
    uint32_t unused;
};
template <bool MULTI>
struct BatchArg {
    using type = NoBatches;
};
template <>
struct BatchArg<true> {
    using type = BatchList;
};
// what the producer tells the consumers about the tile sitting in a stage (MULTI only)
struct StageInfo {
    uint32_t first, rows, raw, pad;
    const uint16_t* in_port;
    uint64_t* result;
};

// ------------------------------------------------------------------ PTX helpers ----------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// same wait, for the producer lane: it only needs to wake up once per tile, so let the hardware park
// it for a while between polls instead of burning issue slots the consumer warps could use
#ifndef SFC_PRODUCER_HINT_NS
#define SFC_PRODUCER_HINT_NS 2000
#endif
#ifndef SFC_PRODUCER_HINT_NS_ACL
#define SFC_PRODUCER_HINT_NS_ACL 16000  // an ACL tile takes several microseconds: poll less often
#endif
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity, uint32_t hint_ns) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAITR_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra DONER_%=;\n\t"
        "bra WAITR_%=;\n\t"
        "DONER_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity), "r"(hint_ns)
        : "memory");
}
__device__ __forceinline__ void tma_load_tile(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar,
                                              uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%2, %3}], [%4], %5;" ::"r"(smem_u32(dst)),
        "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
// plain byte copy global -> shared through the same engine (no tensor map, no swizzle): partial tiles
__device__ __forceinline__ void bulk_load_rows(void* dst, const void* src, uint32_t bytes, uint64_t* bar,
                                               uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
// one key bit of the ACL stage: `field` = row offset | right shift (see sfc_common.h), appended to idx
__device__ __forceinline__ uint32_t acl_key_bit(uint32_t key_base, uint32_t field, uint32_t idx) {
    uint32_t word;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(word) : "r"(key_base + (field & 0xFF80u)));
    return (idx << 1) | ((word >> (field & 31u)) & 1u);
}
// one 32-byte sector, read-only path
__device__ __forceinline__ void ldg256(const void* p, uint32_t* r) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}
// same sector, allowed to stay in L1 (tables small enough to live there)
__device__ __forceinline__ void ldg256_l1(const void* p, uint32_t* r) {
    asm volatile("ld.global.nc.L1::evict_last.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}
// same sector through L1 with the default policy: the ACL rule records (96 B each) are few and hot;
// on B200 letting them live in L1 took configs[3] from 58 to 46 us per Mi packets
__device__ __forceinline__ void ldg256_c(const void* p, uint32_t* r) {
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}
__device__ __forceinline__ void stg_result(uint64_t* p, uint32_t lo, uint32_t hi) {
    asm volatile("st.global.cs.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(lo), "r"(hi) : "memory");
}
__device__ __forceinline__ uint32_t be32_pair(uint32_t lo_word, uint32_t hi_word) {
    // big-endian value of the 4 bytes that start at byte 2 of lo_word
    return __byte_perm(lo_word, hi_word, 0x2345);
}
__device__ __forceinline__ uint32_t word_at(const uint32_t* w, int i) {  // dynamic pick, i in 8..15
    uint32_t v = 0;
#pragma unroll
    for (int k = 8; k < 16; ++k) v = (i == k) ? w[k] : v;
    return v;
}

// ------------------------------------------------------------------ stage 0: parse -------
// w[16]: the record as little-endian words.  Fills k; returns false for non-IP frames.
__device__ __forceinline__ bool parse_header(uint32_t* w, Key& k) {
    uint32_t et = __byte_perm(w[3], 0, 0x4401);
    uint32_t vlan = 0;
    int limit = 64;
    if (et == 0x8100u) {
        vlan = __byte_perm(w[3], 0, 0x4423) & 0xFFFu;  // TCI = bytes 14,15
        // a tag is exactly one word: slide the rest of the frame down by one
#pragma unroll
        for (int i = 3; i < 15; ++i) w[i] = w[i + 1];
        w[15] = 0;
        et = __byte_perm(w[3], 0, 0x4401);
        limit = 60;
    }
    k.meta = vlan << 8;
    k.ip_ver = 0;
    k.ports = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) k.sip[i] = k.dip[i] = 0;
    uint32_t vihl = (w[3] >> 16) & 0xFFu;  // byte 14
    if (et == 0x0800u) {
        uint32_t ihl = vihl & 0xFu;
        if ((vihl >> 4) != 4u || ihl < 5u) return false;
        uint32_t proto = w[5] >> 24;                          // byte 23
        uint32_t frag = __byte_perm(w[5], 0, 0x4401) & 0x1FFFu;  // bytes 20,21
        k.sip[0] = be32_pair(w[6], w[7]);
        k.dip[0] = be32_pair(w[7], w[8]);
        k.meta |= proto;
        k.ip_ver = 4;
        int l4 = 14 + 4 * (int)ihl;
        bool has_ports = (proto == 6u || proto == 17u || proto == 132u) && frag == 0u && l4 + 4 <= limit;
        if (has_ports) {
            if (ihl == 5u)
                k.ports = be32_pair(w[8], w[9]);
            else  // options: L4 starts (ihl-5) words later
                k.ports = be32_pair(word_at(w, 3 + (int)ihl), word_at(w, 4 + (int)ihl));
        }
        return true;
    }
    if (et == 0x86DDu) {
        if ((vihl >> 4) != 6u) return false;
        uint32_t proto = w[5] & 0xFFu;  // byte 20
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            k.sip[i] = be32_pair(w[5 + i], w[6 + i]);
            k.dip[i] = be32_pair(w[9 + i], w[10 + i]);
        }
        k.meta |= proto;
        k.ip_ver = 6;
        if (proto == 6u || proto == 17u || proto == 132u) k.ports = be32_pair(w[13], w[14]);
        return true;
    }
    return false;
}

// ------------------------------------------------------------------ stage 2: flows -------
// One 32-byte sector per probe.  Two strategies, picked per table by the host (flow4_mode):
//  - lazy  (tables that live in HBM): probe the first-choice bucket, which holds most flows (the
//    builder reports how many), and load the second only on a miss — fewest DRAM sectors;
//  - eager (tables that fit L2/L1): issue both loads at once — one latency instead of two.
__device__ __forceinline__ bool flow4_match(const uint32_t* s, const Key& k, uint32_t tag) {
    return s[0] == k.sip[0] && s[1] == k.dip[0] && s[2] == k.ports && s[3] == tag;
}
__device__ __forceinline__ bool flow4_lookup(const DevTables& T, const Key& k, uint32_t& rhi, uint32_t& rlo) {
    const uint32_t meta20 = k.meta & 0xFFFFFu, tag = meta20 | kSlotValid;
    uint32_t x, y;
    flow4_hash(k.sip[0], k.dip[0], k.ports, meta20, T.S4, x, y);
    const uint32_t bw = kFlow4BucketWords * T.flow4_slots;
    const uint32_t* p0 = T.flow4 + (size_t)flow_bucket(x, y, 0, T.S4 + 4, T.flow4_nbuckets) * bw;
    const uint32_t* p1 = T.flow4 + (size_t)flow_bucket(x, y, 1, T.S4 + 4, T.flow4_nbuckets) * bw;
    uint32_t s[8], t[8];
    bool hit;
    if (T.flow4_slots == 2) {
        // dense L2-resident table: the two slots of a bucket are the two halves of one 64-byte line;
        // both are fetched at once, the second bucket only if neither holds the flow
        ldg256(p0, s);
        ldg256(p0 + 8, t);
        hit = flow4_match(s, k, tag);
        if (!hit && flow4_match(t, k, tag)) {
            hit = true;
            s[4] = t[4];
            s[5] = t[5];
        }
        if (!hit) {
            ldg256(p1, s);
            ldg256(p1 + 8, t);
            hit = flow4_match(s, k, tag);
            if (!hit && flow4_match(t, k, tag)) {
                hit = true;
                s[4] = t[4];
                s[5] = t[5];
            }
        }
    } else if (T.flow4_mode & kProbeEager) {
        if (T.flow4_mode & kProbeL1) {
            ldg256_l1(p0, s);
            ldg256_l1(p1, t);
        } else {
            ldg256(p0, s);
            ldg256(p1, t);
        }
        hit = flow4_match(s, k, tag);
        if (!hit) {
            hit = flow4_match(t, k, tag);
            s[4] = t[4];
            s[5] = t[5];
        }
    } else if (T.flow4_mode & kProbeL1) {  // lazy, probes may stay in L1: skewed traffic keeps its hot sectors there
        ldg256_l1(p0, s);
        hit = flow4_match(s, k, tag);
        if (!hit) {
            ldg256_l1(p1, s);
            hit = flow4_match(s, k, tag);
        }
    } else {
        ldg256(p0, s);
        hit = flow4_match(s, k, tag);
        if (!hit) {
            ldg256(p1, s);
            hit = flow4_match(s, k, tag);
        }
    }
    if (!hit && T.flow4_nhash > 2) {
        const uint32_t* p2 = T.flow4 + (size_t)flow_bucket(x, y, 2, T.S4 + 4, T.flow4_nbuckets) * bw;
        for (uint32_t sl = 0; sl < T.flow4_slots && !hit; ++sl) {
            ldg256(p2 + sl * kFlow4BucketWords, s);
            hit = flow4_match(s, k, tag);
        }
    }
    rhi = s[4];
    rlo = s[5];
    return hit;
}
__device__ __forceinline__ bool flow6_probe(const uint32_t* p, const Key& k, uint32_t tag, uint32_t& rhi, uint32_t& rlo, bool l1) {
    uint32_t s[16];
    if (l1) {
        ldg256_l1(p, s);
        ldg256_l1(p + 8, s + 8);
    } else {
        ldg256(p, s);
        ldg256(p + 8, s + 8);
    }
    bool eq = s[8] == k.ports && s[9] == tag;
#pragma unroll
    for (int i = 0; i < 4; ++i) eq = eq && s[i] == k.sip[i] && s[4 + i] == k.dip[i];
    rhi = s[10];
    rlo = s[11];
    return eq;
}
__device__ __forceinline__ bool flow6_lookup(const DevTables& T, const Key& k, uint32_t& rhi, uint32_t& rlo) {
    const uint32_t meta20 = k.meta & 0xFFFFFu, tag = meta20 | kSlotValid;
    uint32_t x, y;
    flow6_hash(k.sip, k.dip, k.ports, meta20, T.S6, x, y);
    bool hit = false;
    for (uint32_t j = 0; j < T.flow6_nhash && !hit; ++j) {
        const uint32_t* p = T.flow6 + (size_t)flow_bucket(x, y, (int)j, T.S6 + 10, T.flow6_nbuckets) * (kFlow6BucketWords * T.flow6_slots);
        const bool l1 = (T.flow4_mode & kProbeL1) != 0;
        hit = flow6_probe(p, k, tag, rhi, rlo, l1);
        if (!hit && T.flow6_slots == 2) hit = flow6_probe(p + kFlow6BucketWords, k, tag, rhi, rlo, l1);
    }
    return hit;
}


// ---- format 2: fingerprint index, then one record (sfc_common.h) ----
// One code path for IPv4 and IPv6: the index probe does not know the family, an IPv6 record is one more
// sector.  A bucket is one sector: seven entries and the count of flows that overflowed into their second
// bucket, so a miss that finds that count at zero stops after one load.
__device__ __forceinline__ bool flow2_probe_unit(const DevTables& T, uint32_t unit, const Key& k, uint32_t tag, bool v6,
                                                 const uint32_t* c, uint32_t& rhi, uint32_t& rlo) {
    const uint32_t* p = T.frec + (size_t)unit * kFrecUnitWords;
    uint32_t u[8], v[8];
    if (T.flow4_mode & kProbeL1) ldg256_l1(p, u); else ldg256(p, u);
    bool eq = u[0] == tag && u[1] == k.ports && u[4] == c[0] && u[5] == c[1] && u[6] == c[2] && u[7] == c[3];
    if (v6) {  // uniform code up to here; the second unit only for IPv6 lanes
        if (T.flow4_mode & kProbeL1) ldg256_l1(p + 8, v); else ldg256(p + 8, v);
        eq = eq && v[0] == k.dip[0] && v[1] == k.dip[1] && v[2] == k.dip[2] && v[3] == k.dip[3];
    }
    rhi = u[2];
    rlo = u[3];
    return eq;
}
__device__ __forceinline__ bool flow2_probe_bucket(const DevTables& T, uint32_t b, uint32_t fpw, uint32_t mask, const Key& k,
                                                   uint32_t tag, bool v6, const uint32_t* c, uint32_t& rhi, uint32_t& rlo,
                                                   uint32_t& overflow) {
    uint32_t e[8];
    const uint32_t* p = T.fidx + (size_t)b * kFidxBucketWords;
    if (T.flow4_mode & kProbeL1) ldg256_l1(p, e); else ldg256(p, e);
    overflow = e[7];
    // first entry carrying our fingerprint, and how many do (more than one is rare: 7 / 2^(32 - bits))
    uint32_t cand = 0, ncand = 0;
#pragma unroll
    for (int i = 0; i < kFidxEntries; ++i) {
        const uint32_t d = e[i] ^ fpw;  // equal fingerprints leave unit + 1 (1 .. mask); an empty entry never passes
        const bool ok = d - 1u < mask;
        cand = (ok && !cand) ? d : cand;
        ncand += ok;
    }
    if (!cand) return false;
    if (flow2_probe_unit(T, cand - 1u, k, tag, v6, c, rhi, rlo)) return true;
    if (ncand > 1) {
        bool skipped = false;
#pragma unroll 1
        for (int i = 0; i < kFidxEntries; ++i) {
            uint32_t ei = 0;
#pragma unroll
            for (int q = 0; q < kFidxEntries; ++q) ei = (q == i) ? e[q] : ei;
            const uint32_t d = ei ^ fpw;
            if (d - 1u >= mask) continue;
            if (!skipped) {  // the first one has been tried
                skipped = true;
                continue;
            }
            if (flow2_probe_unit(T, d - 1u, k, tag, v6, c, rhi, rlo)) return true;
        }
    }
    return false;
}
__device__ __forceinline__ bool flow2_lookup(const DevTables& T, const Key& k, uint32_t& rhi, uint32_t& rlo) {
    const bool v6 = k.ip_ver == 6;
    const uint32_t meta20 = k.meta & 0xFFFFFu, tag = meta20 | kSlotValid | (v6 ? kSlotV6 : 0u);
    uint32_t x, y;
    flow6_hash(k.sip, k.dip, k.ports, flow2_meta(meta20, v6), T.S2, x, y);
    const uint32_t mask = (1u << T.fidx_bits) - 1u;
    const uint32_t fpw = (flow_fp(x, y, T.S2 + 10) >> T.fidx_bits) << T.fidx_bits;
    const uint32_t c[4] = {k.sip[0], v6 ? k.sip[1] : k.dip[0], v6 ? k.sip[2] : 0u, v6 ? k.sip[3] : 0u};
    const uint32_t b0 = flow_bucket(x, y, 0, T.S2 + 10, T.fidx_nbuckets);
    uint32_t overflow;
    if (flow2_probe_bucket(T, b0, fpw, mask, k, tag, v6, c, rhi, rlo, overflow)) return true;
    if (!overflow) return false;
    const uint32_t b1 = flow_bucket(x, y, 1, T.S2 + 10, T.fidx_nbuckets);
    if (b1 == b0) return false;
    uint32_t unused;
    return flow2_probe_bucket(T, b1, fpw, mask, k, tag, v6, c, rhi, rlo, unused);
}

// ------------------------------------------------------------------ stage 4: LPM ---------
__device__ __forceinline__ uint32_t lpm_word_mask(uint32_t len, int i) {  // bits of word i covered by /len
    const int bits = (int)len - 32 * i;
    return bits <= 0 ? 0u : (bits >= 32 ? 0xFFFFFFFFu : ~(0xFFFFFFFFu >> bits));
}
// `a` = the destination as four words; an IPv4 address sits in a[0] with a[1..3] = 0, and its trie never
// has a child below bit 32, so the same walk serves both families.
// Returns the entry value (next-hop index + 1, 0 = no route).  A hit on a tail record also delivers the high
// result word (hi_valid), which saves the next-hop load.
__device__ __forceinline__ uint32_t lpm_lookup(const uint32_t* __restrict__ root, const uint32_t* __restrict__ nodes,
                                               const LpmTail* __restrict__ tails, const uint32_t* a, uint32_t& hi,
                                               bool& hi_valid) {
    hi_valid = false;
    uint32_t e = __ldg(root + (a[0] >> 16));
#pragma unroll
    for (int pos = 16; pos < 128; pos += 8) {  // static positions: the address stays in registers
        if (!(e & kLpmChild)) break;
        uint32_t byte = (a[pos >> 5] >> (24 - (pos & 31))) & 0xFFu;
        e = __ldg(nodes + (size_t)(e & ~kLpmChild) * 256 + byte);
    }
    if (e & kLpmChild) return 0u;
    if (e & kLpmTail) {  // a lone prefix below this point: one record, one masked compare
        uint32_t t[8];
        ldg256(tails + (e & ~kLpmTail), t);
        uint32_t d = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) d |= (a[i] ^ t[i]) & lpm_word_mask(t[4], i);
        e = d ? t[6] : t[5];
        hi = t[7];
        hi_valid = d == 0u;
    }
    return e;
}

// COUNTERS: add a thread's byte-packed tallies (at most 255 each) to the CTA's shared counters, one warp
// reduction and one shared atomic per counter.
__device__ __forceinline__ void fold_counters(unsigned long long* scnt, uint32_t acc_act, unsigned long long acc_stage,
                                              int lane) {
#pragma unroll
    for (uint32_t a = 0; a < 3; ++a) {
        const uint32_t c = __reduce_add_sync(0xFFFFFFFFu, (acc_act >> (8u * a)) & 0xFFu);
        if (lane == 0 && c) atomicAdd(scnt + SFC_CNT_ACTION0 + a, (unsigned long long)c);
    }
#pragma unroll
    for (uint32_t st = 0; st < 7; ++st) {
        const uint32_t c = __reduce_add_sync(0xFFFFFFFFu, (uint32_t)(acc_stage >> (8u * st)) & 0xFFu);
        if (lane == 0 && c) atomicAdd(scnt + SFC_CNT_STAGE0 + st, (unsigned long long)c);
    }
}

// COUNTERS: per-rule hit counters (ACL and L2 rules).  A few rules take most of the hits, and same-address
// atomics from 148 SMs serialise in one L2 slice (27 M of them cost configs[4] 14 ms), so each CTA keeps a small
// direct-mapped cache of counters in shared memory — key = counter index + 1, claimed with a CAS — and only a
// rule whose slot is taken by another rule goes to global memory at once.  The cache is added to the global
// counters when the CTA ends.
constexpr int kRuleCacheSlots = 64;
__device__ __forceinline__ void count_rule(unsigned long long* global, uint32_t* ckey, uint32_t* cval, uint32_t idx) {
    const uint32_t slot = idx & (kRuleCacheSlots - 1), want = idx + 1u;
    uint32_t k = *(volatile uint32_t*)(ckey + slot);
    if (k == 0u) {
        k = atomicCAS(ckey + slot, 0u, want);
        if (k == 0u) k = want;
    }
    if (k == want) atomicAdd(cval + slot, 1u);
    else atomicAdd(global + idx, 1ull);
}

// ------------------------------------------------------------------ the kernel -----------
template <int TILE, int STAGES, bool ACL>
struct SmemLayout {
    static constexpr int kTileBytes = TILE * 64;
    static constexpr int kAclBytes = ACL ? (TILE / 32) * kAclKeyRows * 32 * 4 : 0;
    static constexpr int kBytes = STAGES * kTileBytes + kAclBytes + 2 * STAGES * 8 + 16 * 8 + 32 * 4 + STAGES * 4 + 16 +
                                  STAGES * (int)sizeof(StageInfo) + (TILE / 32) * 8 + 2 * kRuleCacheSlots * 4 + 1024;
};

template <int TILE, int STAGES, bool COUNTERS, bool ACL, bool MULTI = false, int OCC = 0, int FMT = 1>
#ifndef SFC_MINB
#define SFC_MINB 1
#endif
#ifndef SFC_MINB_ACL
#define SFC_MINB_ACL 4  // <= 56 registers: four 288-thread CTAs per SM (cfg4: 67 vs 76 us per Mi packets; 5 would spill)
#endif
// OCC = 5: the same code held to 40 registers so that five CTAs share an SM — slower where the table is in L1
// (more instructions), faster where probes wait on L2 / DRAM (api.cu picks per committed table).
__global__ void __launch_bounds__(TILE + 32, OCC ? OCC : ((ACL || COUNTERS || FMT == 2) ? SFC_MINB_ACL : SFC_MINB)) classify_kernel(const __grid_constant__ CUtensorMap hdr_map,
                                                              const uint16_t* __restrict__ in_port,
                                                              uint64_t* __restrict__ result, uint32_t n,
                                                              uint32_t n_tiles, uint32_t* ticket,
                                                              uint32_t overlap,
                                                              const __grid_constant__ DevTables T,
                                                              const __grid_constant__ typename BatchArg<MULTI>::type L) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* tiles = smem;
    uint32_t* aclq = (uint32_t*)(smem + STAGES * TILE * 64);
    uint64_t* full = (uint64_t*)((uint8_t*)aclq + SmemLayout<TILE, STAGES, ACL>::kAclBytes);
    uint64_t* empty = full + STAGES;
    unsigned long long* scnt = (unsigned long long*)(empty + STAGES);  // 16 block-level counters
    uint32_t* dep_sink = (uint32_t*)(scnt + 16);                       // one word per consumer warp
    uint32_t* tile_of = dep_sink + 32;                                 // tile index held by each stage
    StageInfo* info = (StageInfo*)(((uintptr_t)(tile_of + STAGES) + 15) & ~(uintptr_t)15);
    uint64_t** wres = (uint64_t**)(info + STAGES);  // MULTI: result base of the tile each consumer warp is working on
    uint32_t* ckey = (uint32_t*)(wres + TILE / 32);  // COUNTERS: per-rule counter cache (count_rule)
    uint32_t* cval = ckey + kRuleCacheSlots;

    const int tid = threadIdx.x;
    constexpr int kConsumerWarps = TILE / 32;
    constexpr int kBoxRows = TILE < 256 ? TILE : 256;
    static_assert(TILE % kBoxRows == 0, "TILE must be a multiple of the TMA box height");
    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, kConsumerWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (COUNTERS && tid < 16) scnt[tid] = 0;
    if (COUNTERS && tid < 2 * kRuleCacheSlots) ckey[tid] = 0;  // keys and values are adjacent
    __syncthreads();

    if (tid >= TILE) {
        // ---------------- producer warp: one lane feeds the ring ----------------
        // Tiles are handed out dynamically: the first one is the CTA's own index, the rest come from
        // a global ticket counter, so SMs that run slow (far L2 die, unlucky probes) simply take
        // fewer tiles.  The draw for the next tile is issued before waiting on the current slot.
        // ticket[0] is the draw counter, ticket[1] counts producers that have finished; the last one
        // to finish puts both back to zero, so the slot is ready for its next launch (or graph replay).
        if (tid == TILE) {
            if (overlap) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
            uint64_t pol = policy_evict_first();
            uint32_t tile = blockIdx.x;
            for (uint32_t it = 0;; ++it) {
                const uint32_t s = it % STAGES, use = it / STAGES;
                const bool have = tile < n_tiles;
                uint32_t next = 0;
                if (have) next = gridDim.x + atomicAdd(ticket, 1u);
                if (use > 0) mbar_wait_relaxed(empty + s, (use - 1) & 1, ACL ? SFC_PRODUCER_HINT_NS_ACL : SFC_PRODUCER_HINT_NS);
                tile_of[s] = have ? tile : 0xFFFFFFFFu;
                if (!have) {
                    mbar_arrive(full + s);  // end marker: no bytes will follow
                    break;
                }
                if constexpr (MULTI) {
                    // which batch owns this tile: the last one whose first tile is not beyond it
                    uint32_t lo = 0, hi = L.count;
                    while (hi - lo > 1) {
                        const uint32_t mid = (lo + hi) >> 1;
                        if (L.b[mid].tile0 <= tile) lo = mid; else hi = mid;
                    }
                    const uint32_t first = (tile - L.b[lo].tile0) * TILE, left = L.b[lo].n - first;
                    const uint32_t rows = left < (uint32_t)TILE ? left : (uint32_t)TILE;
                    const uint32_t row = L.b[lo].row0 + first;
                    StageInfo& si = info[s];
                    si.first = first;
                    si.rows = rows;
                    si.raw = rows < (uint32_t)TILE;
                    si.in_port = L.b[lo].in_port;
                    si.result = L.b[lo].result;
                    mbar_arrive_expect_tx(full + s, rows * 64);
                    if (rows == (uint32_t)TILE) {
#pragma unroll
                        for (int r = 0; r < TILE; r += kBoxRows)
                            tma_load_tile(tiles + s * TILE * 64 + r * 64, &hdr_map, 0, (int)row + r, full + s, pol);
                    } else {
                        // the end of a batch: a box would run past it into memory that is not ours, so
                        // copy exactly the rows that exist, unswizzled (consumers are told: raw)
                        bulk_load_rows(tiles + s * TILE * 64, L.base + (size_t)row * 64, rows * 64, full + s, pol);
                    }
                } else {
                    mbar_arrive_expect_tx(full + s, TILE * 64);
#pragma unroll
                    for (int r = 0; r < TILE; r += kBoxRows)  // a TMA box is at most 256 rows
                        tma_load_tile(tiles + s * TILE * 64 + r * 64, &hdr_map, 0, (int)(tile * TILE) + r, full + s, pol);
                }
                tile = next;
            }
            __threadfence();
            if (atomicAdd(ticket + 1, 1u) == gridDim.x - 1) {  // every producer has made its last draw
                ticket[0] = 0;
                ticket[1] = 0;
            }
            // keep stream order intact when launches overlap: do not retire before the previous grid has
            if (overlap) asm volatile("griddepcontrol.wait;" ::: "memory");
        }
        return;
    }
    // ---------------- consumers: one packet per thread per tile ----------------
    const int lane = tid & 31;
    if (ACL) aclq[(tid >> 5) * (kAclKeyRows * 32) + 11 * 32 + lane] = 0u;  // the zero row unused tree bits read
    const uint32_t row_off = (uint32_t)tid * 64u;
    const uint32_t sw = ((uint32_t)tid >> 1) & 3u;
    uint32_t acc_act = 0u, acc_tiles = 0u;  // COUNTERS: byte a = packets of this thread with action a since the last fold
    unsigned long long acc_stage = 0ull;    //           byte s = packets decided at stage s
    for (uint32_t it = 0;; ++it) {
        const uint32_t s = it % STAGES, use = it / STAGES;
        mbar_wait(full + s, use & 1);
        const uint32_t tile = tile_of[s];
        if (tile == 0xFFFFFFFFu) break;
        uint32_t pkt, swz = sw, info_dep = 0;
        bool live;
        const uint16_t* ports_in = in_port;
        uint64_t* res_out = result;
        if constexpr (MULTI) {
            const StageInfo& si = info[s];
            pkt = si.first + tid;
            live = (uint32_t)tid < si.rows;
            if (si.raw) swz = 0;
            ports_in = si.in_port;
            res_out = si.result;
            info_dep = si.first ^ si.rows ^ si.raw ^ (uint32_t)(uintptr_t)ports_in ^ (uint32_t)(uintptr_t)res_out;
            // the result pointer is not needed until the very end of the tile: park it in the warp's own slot
            // instead of carrying it in two registers through the whole lookup
            if (lane == 0) wres[tid >> 5] = res_out;
            __syncwarp();
        } else {
            pkt = tile * TILE + tid;
            live = pkt < n;
        }
        const uint32_t port = live ? (uint32_t)__ldg(ports_in + pkt) : 0u;
        uint32_t w[16];
        const uint32_t base = smem_u32(tiles + s * TILE * 64) + row_off;
#pragma unroll
        for (uint32_t c = 0; c < 4; ++c) {
            uint4 v = lds128(base + ((c ^ swz) << 4));
            w[4 * c + 0] = v.x;
            w[4 * c + 1] = v.y;
            w[4 * c + 2] = v.z;
            w[4 * c + 3] = v.w;
        }
        // Release the slot as soon as the record sits in registers.  "Sits in registers" must be
        // enforced: LDS data returns asynchronously, and under load the arrive below can overtake
        // it, letting the producer's next TMA overwrite rows that have not been read yet (seen on
        // B200 as sporadic wrong rows).  Folding one word of each of the four loads into a value
        // that lane 0 stores makes the store — issued in order before the arrive — wait on the
        // warp-wide scoreboard of all four LDS instructions.
        const uint32_t dep = w[3] ^ w[7] ^ w[11] ^ w[15] ^ tile ^ info_dep;
        __syncwarp();
        if (lane == 0) {
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(smem_u32(dep_sink + (tid >> 5))), "r"(dep) : "memory");
            mbar_arrive(empty + s);
        }

        // dst MAC first: parse_header may slide the words of a tagged frame
        const uint32_t mac_hi = __byte_perm(w[0], 0, 0x4401), mac_lo = be32_pair(w[0], w[1]);
        Key k;
        const bool is_ip = parse_header(w, k);
        uint32_t rhi = kResMissHi, rlo = kResMissLo;
        bool done = !live;  // lanes past the end of the batch hold zero-filled records

        // ---- stage 1: ACL — bit-selecting decision trees, one packet per thread ----
        // Each tree narrows the candidates to a short list in rank order; the first candidate that
        // matches all 11 key words under its mask wins, and the best rank over the trees is the
        // verdict.  The key words sit in shared memory because the tested bit positions are table data.
        if (ACL) {
            const uint32_t qb = smem_u32(aclq + (tid >> 5) * (kAclKeyRows * 32) + lane);
            const uint32_t q0 = (port << 16) | (k.meta >> 8);
            const uint32_t q1 = (k.ip_ver << 24) | ((k.meta & 0xFFu) << 16) | (k.ports >> 16), q2 = k.ports << 16;
            sts32(qb + 0 * 128, q0);
            sts32(qb + 1 * 128, q1);
            sts32(qb + 2 * 128, q2);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                sts32(qb + (3 + i) * 128, k.sip[i]);
                sts32(qb + (7 + i) * 128, k.dip[i]);
            }
            if (live && is_ip) {
                uint32_t best = 0xFFFFFFFFu;
                // full compare of candidate c against the key held in registers; g = its G0 sector
                auto matches = [&](uint32_t c, const uint32_t* g) -> bool {
                    uint32_t d = ((q1 ^ g[0]) & g[4]) | ((q2 ^ g[1]) & g[5]) | ((k.dip[0] ^ g[2]) & g[6]) |
                                 ((k.sip[0] ^ g[3]) & g[7]);
                    if (d) return false;
                    const uint32_t* rec = T.acl_rules + (size_t)(c & kAclRankMask) * kAclRuleWords;
                    uint32_t h[8];
                    if (c & kAclNeedG1) {
                        ldg256_c(rec + 8, h);  // G1: words {0, 8, 9, 10}
                        d = ((q0 ^ h[0]) & h[4]) | ((k.dip[1] ^ h[1]) & h[5]) | ((k.dip[2] ^ h[2]) & h[6]) |
                            ((k.dip[3] ^ h[3]) & h[7]);
                        if (d) return false;
                    }
                    if (c & kAclNeedG2) {
                        ldg256_c(rec + 16, h);  // G2: words {4, 5, 6}
                        d = ((k.sip[1] ^ h[0]) & h[4]) | ((k.sip[2] ^ h[1]) & h[5]) | ((k.sip[3] ^ h[2]) & h[6]);
                    }
                    return d == 0u;
                };
                for (uint32_t t = 0; t < T.acl_ntrees; ++t) {
                    uint4 e = T.acl_roots[t];
                    for (int lvl = 0; lvl < kAclMaxLevels && !(e.x & kAclLeafBit); ++lvl) {  // one 16-byte load per level
                        uint32_t idx = acl_key_bit(qb, e.y & 0xFFFFu, 0u);
                        idx = acl_key_bit(qb, e.y >> 16, idx);
                        idx = acl_key_bit(qb, e.z & 0xFFFFu, idx);
                        idx = acl_key_bit(qb, e.z >> 16, idx);
                        idx = acl_key_bit(qb, e.w & 0xFFFFu, idx);
                        idx = acl_key_bit(qb, e.w >> 16, idx);
                        e = __ldg(T.acl_nodes + (e.x + idx));
                    }
                    if (!(e.x & kAclLeafBit)) continue;  // deeper than any tree the builder makes: a damaged image, not a hang
                    // the first six candidates travel in the entry itself; they are tried two at a time so
                    // that the rule sectors of a pair are in flight together
                    const uint32_t cw[3] = {e.y, e.z, e.w};
                    bool open = true;  // false once this tree cannot improve `best` any more
#pragma unroll
                    for (int p = 0; p < 3; ++p) {
                        if (!open) break;
                        const uint32_t ca = cw[p] & 0xFFFFu, cb = cw[p] >> 16;
                        if (ca == kAclNone || (ca & kAclRankMask) >= best) {
                            open = false;
                            break;
                        }
                        const bool two = cb != kAclNone && (cb & kAclRankMask) < best;
                        uint32_t ga[8], gb[8];
                        ldg256_c(T.acl_rules + (size_t)(ca & kAclRankMask) * kAclRuleWords, ga);
                        if (two) ldg256_c(T.acl_rules + (size_t)(cb & kAclRankMask) * kAclRuleWords, gb);
                        if (matches(ca, ga)) {
                            best = ca & kAclRankMask;
                            open = false;
                        } else if (!two) {
                            open = false;
                        } else if (matches(cb, gb)) {
                            best = cb & kAclRankMask;
                            open = false;
                        }
                    }
                    if (open && (e.x & kAclMoreBit)) {  // rare: more than six candidates and none matched yet
                        const uint16_t* more = T.acl_leaves + (e.x & kAclOffMask);
                        const uint32_t cnt = __ldg(more);
                        for (uint32_t i = 1; i <= cnt; ++i) {
                            const uint32_t c = __ldg(more + i);
                            if ((c & kAclRankMask) >= best) break;
                            uint32_t g[8];
                            ldg256_c(T.acl_rules + (size_t)(c & kAclRankMask) * kAclRuleWords, g);
                            if (matches(c, g)) {
                                best = c & kAclRankMask;
                                break;
                            }
                        }
                    }
                }
                if (best != 0xFFFFFFFFu) {
                    const uint32_t r = __ldg(T.acl_res + best);
                    if (COUNTERS) count_rule(T.counters, ckey, cval, SFC_CNT_RULES0 + (r & 0x0FFFFFFFu));
                    if (r & 0x80000000u) {
                        rhi = res_hi(SFC_ACT_DROP, SFC_NF_NONE, 0);
                        rlo = res_lo(SFC_STAGE_ACL, r);
                        done = true;
                    }
                }
            }
        }

        // ---- stage 2: exact 5-tuple ----
        if (!done && is_ip) {
            uint32_t h = 0, l = 0;
            bool hit = false;
            if constexpr (FMT == (int)kFlowFormatIndexed) {  // the kernel is compiled once per flow table format
                if (T.flow4_count | T.flow6_count) hit = flow2_lookup(T, k, h, l);
            } else {
                if (k.ip_ver == 4) {
                    if (T.flow4_count) hit = flow4_lookup(T, k, h, l);
                } else if (T.flow6_count) {
                    hit = flow6_lookup(T, k, h, l);
                }
            }
            if (hit) {
                rhi = h;
                rlo = l;
                done = true;
            }
        }

        // ---- stage 3: SFC L2 steering ----
        if (!done) {
            if (T.l2x_count) {
                uint32_t key_hi = (port << 16) | mac_hi;
                uint32_t i = l2_hash(key_hi, mac_lo) & T.l2x_mask;
                for (;;) {
                    uint4 e = __ldg(reinterpret_cast<const uint4*>(T.l2x + i));
                    if (!e.w) break;
                    if (e.x == mac_lo && e.y == key_hi) {
                        rlo = e.z;
                        rhi = e.w;
                        done = true;
                        if (COUNTERS) count_rule(T.counters, ckey, cval, SFC_CNT_L2_0 + ((e.z & 0x0FFFFFFFu) % SFC_CNT_L2_N));
                        break;
                    }
                    i = (i + 1) & T.l2x_mask;
                }
            }
            if (!done && T.l2w_count && port < T.l2w_n) {
                uint2 e = __ldg(reinterpret_cast<const uint2*>(T.l2w + port));
                if (e.x) {
                    rhi = e.x;
                    rlo = e.y;
                    done = true;
                    if (COUNTERS) count_rule(T.counters, ckey, cval, SFC_CNT_L2_0 + ((e.y & 0x0FFFFFFFu) % SFC_CNT_L2_N));
                }
            }
        }

        // ---- stage 4: LPM on the destination ----
        // One code path for both families: a warp holding IPv4 and IPv6 packets walks the two tries side
        // by side (max of the depths) instead of one after the other (their sum).
        if (!done && is_ip) {
            const bool v6 = k.ip_ver == 6;
            if (v6 ? T.lpm6_count : T.lpm4_count) {
                uint32_t hi;
                bool hi_valid;
                const uint32_t e = lpm_lookup(v6 ? T.lpm6_root : T.lpm4_root, v6 ? T.lpm6_nodes : T.lpm4_nodes,
                                              v6 ? T.lpm6_tails : T.lpm4_tails, k.dip, hi, hi_valid);
                if (e) {
                    if (hi_valid) {
                        rhi = hi;
                        rlo = res_lo(v6 ? SFC_STAGE_LPM6 : SFC_STAGE_LPM4, e - 1);
                    } else {
                        uint2 r = __ldg(reinterpret_cast<const uint2*>((v6 ? T.lpm6_nh : T.lpm4_nh) + (e - 1)));
                        rhi = r.x;
                        rlo = r.y;
                    }
                }
            }
        }

        if constexpr (MULTI) res_out = wres[tid >> 5];
        if (live) stg_result(res_out + pkt, rlo, rhi);

        if (COUNTERS) {
            // Per-action / per-stage counts ride in three registers, one byte per counter, and are folded
            // into the CTA's shared counters every 255 tiles (a ballot per counter per tile cost 20 % of the
            // kernel on large-table traffic; this costs five instructions).
            if (live) {
                acc_act += 1u << (8u * (rhi >> 24));
                acc_stage += 1ull << (8u * (rlo >> 28));
            }
            if (++acc_tiles == 255u) {
                fold_counters(scnt, acc_act, acc_stage, lane);
                acc_act = 0u;
                acc_stage = 0ull;
                acc_tiles = 0u;
            }
        }
    }
    if (COUNTERS) {
        fold_counters(scnt, acc_act, acc_stage, lane);
        // consumers only (the producer warp has left): named barrier over the TILE consumer threads
        asm volatile("bar.sync 1, %0;" ::"n"(TILE));
        if (tid < 16 && scnt[tid]) atomicAdd(T.counters + tid, scnt[tid]);
        if (tid < kRuleCacheSlots && ckey[tid] && cval[tid]) atomicAdd(T.counters + (ckey[tid] - 1u), (unsigned long long)cval[tid]);
    }
}

}  // namespace sfc
