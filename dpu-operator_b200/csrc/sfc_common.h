// sfc_common.h — arithmetic shared by the host table builder and the device kernels:
// hashes, slot encodings and the table-image header.  Integer-only; every function here is
// compiled for both host and device so the builder places a key exactly where the kernel probes.
#pragma once
#include <stdint.h>

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
    uint32_t mac_hi;  // dst MAC bytes 0..1
    uint32_t mac_lo;  // dst MAC bytes 2..5
};

// ---------------------------------------------------------------- flow table -------------
// Bucketised cuckoo table.  A slot holds the key words followed by two meta words:
//   m0 = proto[0:8) | vlan[8:20) | action[20:22) | id[16:26) << 22
//   m1 = out_port[0:16) | id[0:16) << 16
// action == 3 marks an empty slot (a slot of all-ones bytes is empty).
constexpr uint32_t kSlotEmptyWord = 0xFFFFFFFFu;
constexpr uint32_t kFlowIdBits = 26;

constexpr int kFlow4KeyWords = 3;                      // src, dst, ports
constexpr int kFlow4SlotWords = kFlow4KeyWords + 2;    // 20 bytes
constexpr int kFlow4BucketBytes = 32;                  // one L2 sector
constexpr int kFlow4Slots = kFlow4BucketBytes / (4 * kFlow4SlotWords);  // 1
constexpr int kFlow6KeyWords = 9;                      // src[4], dst[4], ports
constexpr int kFlow6SlotWords = 12;                    // 44 bytes used, padded to 48
constexpr int kFlow6BucketBytes = 64;
constexpr int kFlow6Slots = 1;
constexpr int kMaxHashes = 3;

SFC_HD uint32_t slot_m0(uint32_t meta20, uint32_t action, uint32_t id) {
    return (meta20 & 0xFFFFFu) | ((action & 3u) << 20) | ((id >> 16) << 22);
}
SFC_HD uint32_t slot_m1(uint32_t out_port, uint32_t id) { return (out_port & 0xFFFFu) | (id << 16); }
SFC_HD uint32_t slot_action(uint32_t m0) { return (m0 >> 20) & 3u; }
SFC_HD uint32_t slot_id(uint32_t m0, uint32_t m1) { return ((m0 >> 22) << 16) | (m1 >> 16); }
SFC_HD uint32_t slot_port(uint32_t m1) { return m1 & 0xFFFFu; }

// 64-bit digest of a flow key; the d bucket choices are derived from it.
SFC_HD uint64_t flow4_digest(uint32_t sip, uint32_t dip, uint32_t ports, uint32_t meta20,
                             uint64_t seed) {
    uint64_t k0 = ((uint64_t)sip << 32) | dip;
    uint64_t k1 = ((uint64_t)ports << 32) | (meta20 & 0xFFFFFu);
    return mix64(mix64(k0 ^ seed) + k1);
}
SFC_HD uint64_t flow6_digest(const uint32_t* sip, const uint32_t* dip, uint32_t ports,
                             uint32_t meta20, uint64_t seed) {
    uint64_t h = seed;
    h = mix64(h ^ (((uint64_t)sip[0] << 32) | sip[1]));
    h = mix64(h + (((uint64_t)sip[2] << 32) | sip[3]));
    h = mix64(h ^ (((uint64_t)dip[0] << 32) | dip[1]));
    h = mix64(h + (((uint64_t)dip[2] << 32) | dip[3]));
    h = mix64(h ^ (((uint64_t)ports << 32) | (meta20 & 0xFFFFFu)));
    return h;
}
// j-th bucket choice (j < kMaxHashes) out of nbuckets
SFC_HD uint32_t flow_bucket(uint64_t digest, int j, uint32_t nbuckets) {
    uint32_t h;
    if (j == 0)
        h = (uint32_t)digest;
    else if (j == 1)
        h = (uint32_t)(digest >> 32);
    else
        h = (uint32_t)(mix64(digest ^ 0xA5A5A5A5A5A5A5A5ull) >> 16);
    return mulhi32(h, nbuckets);
}

// ---------------------------------------------------------------- L2 steering ------------
// exact table: open addressing, linear probing, 16-byte entries
struct L2Entry {
    uint32_t key_lo;  // MAC bytes 2..5
    uint32_t key_hi;  // in_port << 16 | MAC bytes 0..1
    uint32_t rule_id;
    uint32_t val;  // bit31 valid | out_port
};
SFC_HD uint32_t l2_hash(uint32_t key_hi, uint32_t key_lo) {
    return (uint32_t)(mix64(((uint64_t)key_hi << 32) | key_lo) >> 32);
}
// wildcard table: indexed by in_port; {val, rule_id}, val bit31 = valid | out_port
struct L2Wild {
    uint32_t val;
    uint32_t rule_id;
};

// ---------------------------------------------------------------- LPM --------------------
// Leaf-pushed multibit trie: root indexed by the top 16 address bits, then 8-bit strides.
// entry: bit31 set -> child node number in bits[0:31); else prefix_id+1 (0 = no prefix).
constexpr uint32_t kLpmChild = 0x80000000u;

// ---------------------------------------------------------------- ACL --------------------
// Canonical key bytes K[0..43]; byte i = (Q[i>>2] >> (24 - 8*(i&3))) & 0xFF with
//   Q0 = in_port<<16 | vlan ; Q1 = ip_ver<<24 | proto<<16 | sport ; Q2 = dport<<16 ;
//   Q3..Q6 = sip[0..3] ; Q7..Q10 = dip[0..3]
constexpr int kAclKeyWords = 11;
constexpr int kAclKeyBytes = 44;
constexpr int kAclMaxPos = 44;

// ---------------------------------------------------------------- image ------------------
constexpr uint64_t kImageMagic = 0x3130434653303042ull;  // "B00SFC01"
constexpr uint32_t kImageVersion = 3;
constexpr size_t kImageAlign = 256;

struct ImageHeader {
    uint64_t magic;
    uint32_t version;
    uint32_t header_bytes;
    uint64_t total_bytes;
    uint64_t epoch;

    uint64_t flow4_off;
    uint64_t flow4_seed;
    uint32_t flow4_nbuckets, flow4_nhash, flow4_count, flow4_pad;

    uint64_t flow6_off;
    uint64_t flow6_seed;
    uint32_t flow6_nbuckets, flow6_nhash, flow6_count, flow6_pad;

    uint64_t l2x_off;
    uint32_t l2x_mask, l2x_count;
    uint64_t l2w_off;
    uint32_t l2w_n, l2w_count;
    uint64_t nf_off;
    uint32_t nf_n, nf_count;

    uint64_t lpm4_root_off, lpm4_nodes_off, lpm4_nh_off;
    uint32_t lpm4_nnodes, lpm4_count;
    uint64_t lpm6_root_off, lpm6_nodes_off, lpm6_nh_off;
    uint32_t lpm6_nnodes, lpm6_count;

    uint64_t acl_bv_off, acl_base_off, acl_rule_off;
    uint32_t acl_nrules, acl_nwords;
    uint32_t acl_npos4, acl_npos6;
    uint8_t acl_pos4[kAclMaxPos];  // key byte index of each active position for v4 packets
    uint8_t acl_pos6[kAclMaxPos];
    uint8_t acl_slot4[kAclMaxPos];  // which stored position table (0..acl_nstored) it uses
    uint8_t acl_slot6[kAclMaxPos];
    uint32_t acl_nstored, acl_pad;
};
static_assert(sizeof(ImageHeader) <= 512, "image header grew past its reserved 512 bytes");
constexpr size_t kImageHeaderBytes = 512;

}  // namespace sfc
