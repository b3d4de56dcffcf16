// tables.h — host-side rule store and table-image builder (no CUDA in this file).
//
// The rule store mirrors what the reference's dataplane backends keep implicitly inside OVS:
// ports by name (ovs-dp/ovsdp.go:40-62), steering flows by match (ovs-dp/ovsdp.go:113-162),
// plus the synthesized tables of SURVEY.md §8a a12.  build_image() turns the staged rule set
// into one contiguous, position-independent blob (ImageHeader + sections) that is uploaded to
// HBM as is and can be broadcast rank to rank.
#pragma once
#include <array>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/b200sfc.h"
#include "sfc_common.h"

namespace sfc {

struct FlowRec {
    uint32_t sip[4], dip[4];
    uint32_t ports, meta;  // meta = proto | vlan << 8
    uint32_t id;
    uint8_t ip_ver, action;
    uint16_t out_port;
};

struct L2Rule {
    uint16_t in_port, out_port;
    bool has_mac;
    uint8_t mac[6];
    uint32_t priority;  // OpenFlow priority: 10 / 100 / 32768 (ovs-dp/ovsdp.go:120,127,135)
    uint32_t id;
};

struct LpmRec {
    uint32_t w[4];
    int len;
    uint16_t out_port;
    uint32_t id;
};

struct AclRec {
    sfc_acl_rule r;
    uint32_t id;
};

// open-addressing index over FlowRec (40-byte keys), for duplicate detection and delete
class FlowIndex {
  public:
    void clear();
    // returns position in recs or -1
    int64_t find(const std::vector<FlowRec>& recs, const FlowRec& k) const;
    void insert(const std::vector<FlowRec>& recs, uint32_t pos);
    void rebuild(const std::vector<FlowRec>& recs, size_t expect);
    // Remove recs[pos] from BOTH the index and the vector in O(1): the last record moves into the
    // hole and its index slot is re-pointed (backward-shift deletion keeps probe chains intact).
    void erase(std::vector<FlowRec>& recs, uint32_t pos);

  private:
    std::vector<uint32_t> slots_;  // pos+1, 0 = empty
    size_t used_ = 0;
    static uint64_t hash(const FlowRec& k);
    void grow(const std::vector<FlowRec>& recs, size_t want);
    int64_t slot_of(const std::vector<FlowRec>& recs, uint32_t pos) const;
};

// One cuckoo flow table (IPv4: 8-word buckets, IPv6: 16-word buckets) as plain host memory, with
// the insert / erase walks.  Used twice: by build_image() on the image being assembled, and as the
// host SHADOW of the committed table so that flow adds and deletes can be applied bucket by bucket
// (sfc_commit's incremental path) instead of rebuilding and re-uploading the whole image.
struct FlowTable {
    bool v6 = false;
    uint32_t* tab = nullptr;  // nbuckets * slots slots of words() u32; slot s of bucket b = b * slots + s
    uint32_t nbuckets = 0;
    int slots = 1;
    int nhash = 2;
    uint32_t S[16] = {0};
    uint64_t rng = 0x1234567ull;

    int words() const { return v6 ? kFlow6BucketWords : kFlow4BucketWords; }
    int tag_index() const { return v6 ? 9 : 3; }
    void buckets_of(const uint32_t* rec, uint32_t* b) const;
    // rec = one slot image (see encode()); the index of every slot written is appended to *dirty if given
    bool insert(const uint32_t* rec, std::vector<uint32_t>* dirty);
    bool erase(const uint32_t* rec, std::vector<uint32_t>* dirty);
    uint32_t count_first() const;
    static void encode(const FlowRec& f, uint32_t next_nf, uint32_t* rec);  // fills words() words
};

// The indexed flow table (format 2, see sfc_common.h) as plain host memory: a fingerprint index in front of
// a dense record array, both families in one structure.  Like FlowTable it is used on the image being
// assembled and as the host shadow of the committed image; every sector it writes is reported (as a
// sector number relative to the start of the image) so that a commit can ship exactly those.
struct IndexedFlows {
    uint32_t* idx = nullptr;  // nbuckets buckets of kFidxBucketWords words
    uint32_t* rec = nullptr;  // units + 1 units of kFrecUnitWords words
    uint32_t nbuckets = 0, bits = 0, units = 0, used = 0;
    uint32_t idx_sector0 = 0, rec_sector0 = 0;  // image sector numbers of idx[0] and rec[0]
    uint32_t S[kFlow6SeedWords] = {0};
    uint64_t rng = 0x7654321ull;
    std::vector<uint32_t> free1, free2;  // released single units / unit pairs (first unit of the pair)
    uint32_t count = 0, second = 0;      // flows stored; flows living in their second bucket

    struct Where {
        uint32_t b0, b1, fp;
    };
    Where where(const uint32_t* sip, const uint32_t* dip, uint32_t ports, uint32_t meta20, bool v6) const;
    Where where_of_unit(uint32_t unit) const;  // recomputed from the record the unit holds
    // returns the unit of the flow or -1; *bucket / *slot receive its index position
    int64_t find(const FlowRec& f, uint32_t* bucket = nullptr, int* slot = nullptr) const;
    bool insert(const FlowRec& f, uint32_t next_nf, std::vector<uint32_t>* dirty);
    bool erase(const FlowRec& f, std::vector<uint32_t>* dirty);
    long room() const;  // flows that can still be added before the table has to be rebuilt larger (approximate)

  private:
    bool place(uint32_t unit, Where w, std::vector<uint32_t>* dirty);
    void touch_idx(uint32_t bucket, std::vector<uint32_t>* dirty) const {
        if (dirty) dirty->push_back(idx_sector0 + bucket);
    }
    void touch_rec(uint32_t unit, std::vector<uint32_t>* dirty) const {
        if (dirty) dirty->push_back(rec_sector0 + unit);
    }
};

struct FlowChange {
    bool add;
    FlowRec rec;
};

struct RuleStore {
    // ports
    std::map<std::string, uint16_t> port_by_name;
    std::vector<std::string> port_names;  // index = id, "" = free
    // L2 steering
    std::vector<L2Rule> l2;
    uint32_t next_l2_id = 0;
    // chain
    std::vector<uint16_t> nf_in, nf_out;
    // flows
    std::vector<FlowRec> flows;
    FlowIndex flow_index;
    uint32_t next_flow_id = 0;
    std::vector<FlowChange> flow_log;  // flow edits since the last commit, in order
    bool l2_dirty = false;             // L2 steering rules changed since the last commit (rebuilt into the idle copy)
    bool other_dirty = true;           // anything else changed since the last commit: the next commit rebuilds the image
    // LPM
    typedef std::array<uint32_t, 5> LpmKey;  // len, w[0..3]
    std::map<LpmKey, LpmRec> lpm4, lpm6;
    uint32_t next_lpm_id = 0;
    // ACL
    std::vector<AclRec> acl;
    uint32_t next_acl_id = 0;
    // knobs
    int flow_load_pct = 0;  // 0 = the format's default (sector: 40, indexed: 65)
    int flow_hashes = 2;
    int flow_slots = 0;  // slots per flow bucket: 2 = dense two-slot buckets (see build_cuckoo), anything else = 1
    int flow_format = 0;  // 0 = by size, kFlowFormatSector, kFlowFormatIndexed

    void clear_rules();
};

// Converts a public flow key to the internal record; returns false if the key is malformed.
bool flow_from_key(const sfc_flow_key& k, const sfc_flow_action& a, FlowRec& out, std::string& err);

// Build the image.  Returns false and sets err on failure (e.g. cuckoo build could not converge).
bool build_image(const RuleStore& rs, uint64_t epoch, std::vector<uint8_t>& image, std::string& err);
// Rebuild the L2 steering tables of copy `copy` (0 / 1) inside an existing image from the rule store and make
// it the active one in the header.  Returns false if the rules no longer fit the capacity the image was built
// with (the caller then rebuilds the image).  Sectors whose bytes changed are appended to *dirty.
bool rebuild_l2_copy(const RuleStore& rs, uint8_t* image, int copy, std::vector<uint32_t>* dirty);
uint8_t nf_of_port(const RuleStore& rs, uint32_t port);

// Structural validation of an image received from elsewhere (import / broadcast).
// Returns nullptr if the image is sound, else a static message.  Touches no heap.
const char* validate_image(const uint8_t* img, size_t len);

}  // namespace sfc
