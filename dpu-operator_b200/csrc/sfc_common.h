// sfc_common.h — arithmetic shared by the host table builder and the device kernels:
// hashes, slot encodings and the table-image header.  Integer-only; every function here is
// compiled for both host and device so the builder places a key exactly where the kernel probes.
#pragma once
#include <stdint.h>

#include "../../include/b200sfc.h"

#if defined(__CUDACC__)
#define SFC_HD __host__ __device__ __forceinline__
#else
#define SFC_HD inline
#endif

namespace sfc {

constexpr uint64_t kGamma = 0x9E3779B97F4A7C15ull;

// splitmix64 finaliser
SFC_HD uint64_t mix64(uint64_t z) {
    z ^= z >> 30;
    z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27;
    z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}
// k-th output of splitmix64 started at `seed` (counter based: any k in O(1))
SFC_HD uint64_t draw(uint64_t seed, uint64_t k) { return mix64(seed + (k + 1) * kGamma); }

SFC_HD uint32_t mulhi32(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }

// ---------------------------------------------------------------- canonical parsed key ----
// sip/dip: address words, big-endian value of each 4-byte group (IPv4 uses word 0 only).
// ports = sport<<16 | dport ; meta = proto | vlan<<8 (20 bits)
struct Key {
    uint32_t sip[4];
    uint32_t dip[4];
    uint32_t ports;
    uint32_t meta;
    uint32_t ip_ver;  // 0 (non-IP), 4, 6
};

// ---------------------------------------------------------------- result words -----------
// Every table stores the two halves of the result word it yields, ready to be written out:
//   hi = action<<24 | next_nf<<16 | out_port ;  lo = stage<<28 | index
SFC_HD uint32_t res_hi(uint32_t action, uint32_t nf, uint32_t port) { return (action << 24) | (nf << 16) | port; }
SFC_HD uint32_t res_lo(uint32_t stage, uint32_t index) { return (stage << 28) | (index & 0x0FFFFFFFu); }
constexpr uint32_t kResMissHi = (SFC_ACT_TO_CP << 24) | (SFC_NF_NONE << 16);
constexpr uint32_t kResMissLo = 0;

// ---------------------------------------------------------------- flow table -------------
// Cuckoo table, d = 2 or 3 bucket choices, one slot per bucket (two for tables that only fit in L2 when
// packed, see build_cuckoo).  An IPv4 slot is exactly one 32-byte sector (one LDG.256, one L2/DRAM sector
// per probe); an IPv6 slot is 64 bytes.
//   v4 slot words: [sip, dip, ports, tag, res_hi, res_lo, 0, 0]
//   v6 slot words: [sip0..3, dip0..3, ports, tag, res_hi, res_lo, 0, 0, 0, 0]
// tag = proto | vlan<<8 | 0x80000000 (the top bit marks an occupied slot; empty slots are zero).
constexpr uint32_t kSlotValid = 0x80000000u;
constexpr int kFlow4BucketWords = 8;
constexpr int kFlow6BucketWords = 16;
constexpr int kFlow4BucketBytes = 32;
constexpr int kFlow6BucketBytes = 64;
constexpr int kMaxHashes = 3;
constexpr uint64_t kFlowL2ResidentBytes = 48ull << 20;  // measured on B200: a table up to ~42 MB stays in L2 beside the stream
constexpr uint64_t kFlowDenseLoadPct = 80;              // two-slot buckets, two choices: threshold ~0.89
constexpr int kFlow4SeedWords = 10;  // 4 for the fold, 2 per bucket choice
constexpr int kFlow6SeedWords = 16;  // 10 for the fold, 2 per bucket choice

// Hashing is built from 32x32->64 multiplies folded by XOR (one IMAD.WIDE + one LOP3 each on the
// GPU).  Table correctness never depends on hash quality — every probe compares the full key —
// only build convergence and probe counts do; the builder re-seeds if an insertion walk fails.
SFC_HD uint64_t mul_wide(uint32_t a, uint32_t b) { return (uint64_t)a * b; }

// two 32-bit hashes (+ a third on demand) of an IPv4 5-tuple; S = kFlow4SeedWords seed words
SFC_HD void flow4_hash(uint32_t sip, uint32_t dip, uint32_t ports, uint32_t meta20, const uint32_t* S,
                       uint32_t& x, uint32_t& y) {
    uint64_t P = mul_wide(sip ^ S[0], dip ^ S[1]);
    uint64_t Q = mul_wide(ports ^ S[2], meta20 ^ S[3]);
    x = (uint32_t)P ^ (uint32_t)(Q >> 32);
    y = (uint32_t)(P >> 32) ^ (uint32_t)Q;
}
SFC_HD void flow6_hash(const uint32_t* sip, const uint32_t* dip, uint32_t ports, uint32_t meta20, const uint32_t* S,
                       uint32_t& x, uint32_t& y) {
    uint64_t P1 = mul_wide(sip[0] ^ S[0], sip[1] ^ S[1]);
    uint64_t P2 = mul_wide(sip[2] ^ S[2], sip[3] ^ S[3]);
    uint64_t P3 = mul_wide(dip[0] ^ S[4], dip[1] ^ S[5]);
    uint64_t P4 = mul_wide(dip[2] ^ S[6], dip[3] ^ S[7]);
    uint64_t Q = mul_wide(ports ^ S[8], meta20 ^ S[9]);
    x = (uint32_t)P1 ^ (uint32_t)(P2 >> 32) ^ (uint32_t)P3 ^ (uint32_t)(P4 >> 32) ^ (uint32_t)(Q >> 32);
    y = (uint32_t)(P1 >> 32) ^ (uint32_t)P2 ^ (uint32_t)(P3 >> 32) ^ (uint32_t)P4 ^ (uint32_t)Q;
}
// j-th bucket choice from the folded pair (x, y); T = 6 seed words (two per choice)
SFC_HD uint32_t flow_bucket(uint32_t x, uint32_t y, int j, const uint32_t* T, uint32_t nbuckets) {
    uint64_t F = mul_wide(x ^ T[2 * j], y ^ T[2 * j + 1]);
    return mulhi32((uint32_t)F ^ (uint32_t)(F >> 32), nbuckets);
}
// expands a 64-bit seed into n odd-ish 32-bit seed words
SFC_HD void expand_seed(uint64_t seed, uint32_t* out, int n) {
    for (int i = 0; i < n; ++i) out[i] = (uint32_t)(draw(seed, (uint64_t)i) >> 16) | 1u;
}

// ---------------------------------------------------------------- flow table, format 2 ----
// "Indexed" format: one fingerprint index for both families in front of a dense record array, so that
// what a lookup touches is small enough to stay in L2 (and, for small rule sets, in L1):
//   index   buckets of one 32-byte sector: 7 entries + 1 overflow word.  entry = 0 (empty) or
//           fp << B | (unit + 1), B = header.fidx_bits, fp = the top 32 - B bits of flow_fp().  Two
//           bucket choices; a flow sits in its first choice unless that was full.  Word 7 counts the
//           flows whose first choice is this bucket but which live in their second one: a lookup
//           goes on to the second bucket only when that count is not zero (misses cost one sector).
//   records 32-byte units.  IPv4: one unit [tag, ports, res_hi, res_lo, sip, dip, 0, 0].  IPv6: two
//           consecutive units [tag, ports, res_hi, res_lo, sip0..3] [dip0..3, 0, 0, 0, 0].
//           tag = proto | vlan << 8 | kSlotV6 (IPv6) | kSlotValid.
// The hash is flow6_hash over the full key (IPv4 addresses in word 0, the rest zero) for both families.
constexpr uint32_t kSlotV6 = 0x40000000u;
constexpr int kFidxEntries = 7;
constexpr int kFidxBucketWords = 8;
constexpr int kFrecUnitWords = 8;
constexpr uint32_t kFlowFormatSector = 1, kFlowFormatIndexed = 2;
SFC_HD uint32_t flow_fp(uint32_t x, uint32_t y, const uint32_t* T) {  // T = the six bucket-choice seed words
    uint64_t F = mul_wide(x ^ T[4], y ^ T[5]);
    return (uint32_t)F ^ (uint32_t)(F >> 32);
}
// meta word fed to the hash: proto | vlan << 8 | (IPv6 ? 1 << 20 : 0)
SFC_HD uint32_t flow2_meta(uint32_t meta20, bool v6) { return meta20 | (v6 ? 0x100000u : 0u); }

// ---------------------------------------------------------------- L2 steering ------------
// exact table: open addressing, linear probing, 16-byte entries; res_hi == 0 marks an empty entry
struct L2Entry {
    uint32_t key_lo;  // MAC bytes 2..5
    uint32_t key_hi;  // in_port << 16 | MAC bytes 0..1
    uint32_t res_lo;
    uint32_t res_hi;
};
SFC_HD uint32_t l2_hash(uint32_t key_hi, uint32_t key_lo) {
    uint64_t P = mul_wide(key_hi ^ 0x9E3779B1u, key_lo ^ 0x85EBCA6Bu);
    return (uint32_t)P ^ (uint32_t)(P >> 32);
}
// wildcard table: indexed by in_port; res_hi == 0 marks "no rule"
struct L2Wild {
    uint32_t res_hi;
    uint32_t res_lo;
};
// LPM next hops: indexed by prefix id
struct LpmRes {
    uint32_t res_hi;
    uint32_t res_lo;
};

// ---------------------------------------------------------------- LPM --------------------
// Leaf-pushed multibit trie: root indexed by the top 16 address bits, then 8-bit strides.
// entry: bit31 set -> child node number in bits[0:31); else prefix_id+1 (0 = no prefix).
constexpr uint32_t kLpmChild = 0x80000000u;
// A sub-trie that holds a single prefix is not expanded into 256-entry nodes: the slot points at one
// 32-byte record instead, and the lookup ends there with one masked compare.
constexpr uint32_t kLpmTail = 0x40000000u;
struct LpmTail {
    uint32_t w[4];      // the prefix, masked to its length
    uint32_t len;       // 1..128
    uint32_t match;     // entry value (next-hop index + 1) if the address is inside the prefix
    uint32_t fallback;  // entry value of the covering shorter prefix otherwise (0 = none)
    uint32_t match_hi;  // high result word of `match`, so that a hit needs no next-hop load
};
static_assert(sizeof(LpmTail) == 32, "one sector");

// ---------------------------------------------------------------- ACL --------------------
// Canonical key bytes K[0..43]; byte i = (Q[i>>2] >> (24 - 8*(i&3))) & 0xFF with
//   Q0 = in_port<<16 | vlan ; Q1 = ip_ver<<24 | proto<<16 | sport ; Q2 = dport<<16 ;
//   Q3..Q6 = sip[0..3] ; Q7..Q10 = dip[0..3]
constexpr int kAclKeyWords = 11;
constexpr int kAclKeyBytes = 44;
constexpr int kAclKeyBits = 352;
// Decision-tree entry: 16 bytes (one LDG.128).  An entry describes the CHILD it leads to, so a level
// costs one load and the leaf list needs none for its first six candidates.
//   inner  x = offset of the child's entry array (bit 31 clear); the child tests up to six key bits and
//            has 2^bits entries; y, z, w hold six 16-bit fields, first = most significant index bit:
//            field = key_word * 128 | right_shift  (128 = one row of the per-warp key array in shared
//            memory, right_shift = 31 - bit position inside the word).  Unused fields name row 11,
//            which holds zero, and come first — so the index stays below 2^bits without a branch.
//   leaf   x = kAclLeafBit | (kAclMoreBit | offset into the overflow list, if more than six candidates);
//            y, z, w = six u16 candidates rank | needG1 << 12 | needG2 << 13, ranks ascending,
//            0xFFFF = none.
// Overflow list (u16): count, then the candidates beyond the sixth.
// Rule record (96 B, three 32-byte groups so the common case is ONE sector):
//   G0 = value, mask of key words {1, 2, 7, 3}   (ip_ver/proto/sport, dport, dst[0], src[0])
//   G1 = value, mask of key words {0, 8, 9, 10}  (in_port/vlan, dst[1..3])
//   G2 = value of words {4, 5, 6}, pad, mask of words {4, 5, 6}, pad   (src[1..3])
struct AclEntry {
    uint32_t x, y, z, w;
};
constexpr int kAclMaxNodeBits = 6;
constexpr int kAclInline = 6;
constexpr int kAclKeyRows = 12;  // 11 key words + the zero row
constexpr uint32_t kAclZeroField = 11u * 128u;
constexpr int kAclRuleWords = 24;
constexpr int kAclMaxTrees = 8;
constexpr int kAclMaxLevels = 16;  // the builder stops at depth 12; the kernel never walks further than this
constexpr uint32_t kAclLeafBit = 0x80000000u, kAclMoreBit = 0x40000000u, kAclOffMask = 0x0FFFFFFFu;
constexpr uint16_t kAclNeedG1 = 1u << 12, kAclNeedG2 = 1u << 13, kAclRankMask = 0x0FFFu, kAclNone = 0xFFFFu;

// ---------------------------------------------------------------- image ------------------
constexpr uint64_t kImageMagic = 0x3130434653303042ull;  // "B00SFC01"
constexpr uint32_t kImageVersion = 12;
constexpr size_t kImageAlign = 256;

struct ImageHeader {
    uint64_t magic;
    uint32_t version;
    uint32_t header_bytes;
    uint64_t total_bytes;
    uint64_t epoch;

    uint64_t flow4_off;
    uint64_t flow4_seed;
    uint32_t flow4_nbuckets, flow4_nhash, flow4_count, flow4_first;  // first: slots found by probe 0

    uint64_t flow6_off;
    uint64_t flow6_seed;
    uint32_t flow6_nbuckets, flow6_nhash, flow6_count, flow6_first;

    uint64_t l2x_off;
    uint32_t l2x_mask, l2x_count;
    uint64_t l2w_off;
    uint32_t l2w_n, l2w_count;

    uint64_t lpm4_root_off, lpm4_nodes_off, lpm4_nh_off;
    uint32_t lpm4_nnodes, lpm4_count;
    uint64_t lpm6_root_off, lpm6_nodes_off, lpm6_nh_off;
    uint32_t lpm6_nnodes, lpm6_count;

    // ACL: a forest of bit-selecting decision trees over the 352-bit key, leaves scanned in rank order
    uint64_t acl_nodes_off, acl_leaf_off, acl_rule_off, acl_res_off;
    uint32_t acl_nrules, acl_nentries, acl_nleafwords, acl_ntrees;
    AclEntry acl_roots[kAclMaxTrees];

    uint64_t lpm4_tail_off, lpm6_tail_off;  // LpmTail records
    uint32_t lpm4_ntails, lpm6_ntails;

    // slots per flow bucket: 1 (one sector per probe: tiny and huge tables) or 2 (dense tables that fit
    // in L2 only when packed); a table is nbuckets * slots slots, slot s of bucket b at index b * slots + s
    uint32_t flow4_slots, flow6_slots;

    // flow table format: kFlowFormatSector (the two per-family cuckoo tables above) or kFlowFormatIndexed
    // (fidx_* / frec_*; the flow4_* / flow6_* sections are then empty apart from the counts)
    uint32_t flow_format, fidx_bits;
    uint64_t fidx_off, fidx_seed, frec_off;
    uint32_t fidx_nbuckets, frec_units;  // frec_units: capacity of the record array in 32-byte units (+1 pad unit follows)
    uint32_t frec_used, fidx_first;      // units handed out so far (high-water mark); flows sitting in their first bucket

    // L2 steering tables exist twice (copy 0 / copy 1) so that a commit can rebuild the idle copy while
    // launches read the active one; l2x_off / l2w_off above always describe the ACTIVE copy
    uint64_t l2x_copy_off[2], l2w_copy_off[2];
    uint32_t l2x_cap, l2_active;
};
static_assert(sizeof(ImageHeader) <= 512, "image header grew past its reserved 512 bytes");
constexpr size_t kImageHeaderBytes = 512;

}  // namespace sfc
