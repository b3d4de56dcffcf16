// tables.cpp — host-side table-image builder.  See tables.h.
#include "tables.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace sfc {

// ------------------------------------------------------------------ FlowIndex ------------
uint64_t FlowIndex::hash(const FlowRec& k) {
    uint64_t h = 0x243F6A8885A308D3ull ^ k.ip_ver;
    for (int i = 0; i < 4; ++i) h = mix64(h ^ (((uint64_t)k.sip[i] << 32) | k.dip[i]));
    return mix64(h ^ (((uint64_t)k.ports << 32) | k.meta));
}
static bool same_key(const FlowRec& a, const FlowRec& b) {
    return a.ip_ver == b.ip_ver && a.ports == b.ports && a.meta == b.meta &&
           !memcmp(a.sip, b.sip, 16) && !memcmp(a.dip, b.dip, 16);
}
void FlowIndex::clear() {
    slots_.clear();
    used_ = 0;
}
int64_t FlowIndex::find(const std::vector<FlowRec>& recs, const FlowRec& k) const {
    if (slots_.empty()) return -1;
    size_t mask = slots_.size() - 1;
    for (size_t i = hash(k) & mask;; i = (i + 1) & mask) {
        uint32_t s = slots_[i];
        if (!s) return -1;
        if (same_key(recs[s - 1], k)) return (int64_t)s - 1;
    }
}
void FlowIndex::grow(const std::vector<FlowRec>& recs, size_t want) {
    size_t cap = 1024;
    while (cap < want * 2) cap <<= 1;
    slots_.assign(cap, 0);
    used_ = 0;
    for (uint32_t p = 0; p < recs.size(); ++p) {
        size_t mask = cap - 1, i = hash(recs[p]) & mask;
        while (slots_[i]) i = (i + 1) & mask;
        slots_[i] = p + 1;
        ++used_;
    }
}
void FlowIndex::rebuild(const std::vector<FlowRec>& recs, size_t expect) {
    grow(recs, std::max(expect, recs.size()));
}
void FlowIndex::insert(const std::vector<FlowRec>& recs, uint32_t pos) {
    if (slots_.empty() || (used_ + 1) * 2 > slots_.size()) {
        // recs already contains pos; grow() re-inserts everything including it
        grow(recs, std::max<size_t>(recs.size() * 2, 512));
        return;
    }
    size_t mask = slots_.size() - 1, i = hash(recs[pos]) & mask;
    while (slots_[i]) i = (i + 1) & mask;
    slots_[i] = pos + 1;
    ++used_;
}

int64_t FlowIndex::slot_of(const std::vector<FlowRec>& recs, uint32_t pos) const {
    size_t mask = slots_.size() - 1;
    for (size_t i = hash(recs[pos]) & mask;; i = (i + 1) & mask) {
        if (!slots_[i]) return -1;
        if (slots_[i] == pos + 1) return (int64_t)i;
    }
}
void FlowIndex::erase(std::vector<FlowRec>& recs, uint32_t pos) {
    size_t mask = slots_.size() - 1;
    int64_t hole = slot_of(recs, pos);
    uint32_t last = (uint32_t)recs.size() - 1;
    int64_t last_slot = last != pos ? slot_of(recs, last) : -1;
    if (last_slot >= 0) slots_[last_slot] = pos + 1;  // the last record is about to move to `pos`
    // backward-shift deletion starting at the hole
    size_t i = (size_t)hole;
    slots_[i] = 0;
    for (size_t j = (i + 1) & mask; slots_[j]; j = (j + 1) & mask) {
        uint32_t p = slots_[j] - 1;
        const FlowRec& r = (p == pos && last != pos) ? recs[last] : recs[p];  // slot re-pointed above still hashes as `last`
        size_t home = hash(r) & mask;
        bool movable = ((j - home) & mask) >= ((j - i) & mask);
        if (movable) {
            slots_[i] = slots_[j];
            slots_[j] = 0;
            i = j;
        }
    }
    --used_;
    if (last != pos) recs[pos] = recs[last];
    recs.pop_back();
}

void RuleStore::clear_rules() {
    l2.clear();
    next_l2_id = 0;
    nf_in.clear();
    nf_out.clear();
    flows.clear();
    flow_index.clear();
    next_flow_id = 0;
    flow_log.clear();
    other_dirty = true;
    lpm4.clear();
    lpm6.clear();
    next_lpm_id = 0;
    acl.clear();
    next_acl_id = 0;
}

static uint32_t be32(const uint8_t* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}

bool flow_from_key(const sfc_flow_key& k, const sfc_flow_action& a, FlowRec& out, std::string& err) {
    if (k.ip_ver != 4 && k.ip_ver != 6) {
        err = "flow key: ip_ver must be 4 or 6";
        return false;
    }
    if (k.vlan > 4095) {
        err = "flow key: vlan must be 0..4095";
        return false;
    }
    if (a.action > SFC_ACT_DROP) {
        err = "flow action: unknown action";
        return false;
    }
    if (a.action != SFC_ACT_FORWARD && a.out_port != 0) {
        err = "flow action: out_port must be 0 unless action is FORWARD";
        return false;
    }
    if (a.out_port >= SFC_MAX_PORTS) {
        err = "flow action: out_port out of range";
        return false;
    }
    if (k.ip_ver == 4)
        for (int i = 4; i < 16; ++i)
            if (k.src[i] || k.dst[i]) {
                err = "flow key: IPv4 address bytes 4..15 must be zero";
                return false;
            }
    for (int i = 0; i < 4; ++i) {
        out.sip[i] = be32(k.src + 4 * i);
        out.dip[i] = be32(k.dst + 4 * i);
    }
    out.ports = ((uint32_t)k.sport << 16) | k.dport;
    out.meta = (uint32_t)k.proto | ((uint32_t)k.vlan << 8);
    out.ip_ver = k.ip_ver;
    out.action = a.action;
    out.out_port = a.out_port;
    out.id = 0;
    return true;
}

// ------------------------------------------------------------------ FlowTable ------------
uint8_t nf_of_port(const RuleStore& rs, uint32_t port) {
    for (size_t i = 0; i < rs.nf_in.size(); ++i)
        if (rs.nf_in[i] == port || rs.nf_out[i] == port) return (uint8_t)i;  // lowest NF index wins
    return (uint8_t)SFC_NF_NONE;
}

void FlowTable::encode(const FlowRec& f, uint32_t next_nf, uint32_t* rec) {
    const bool six = f.ip_ver == 6;
    const int bw = six ? kFlow6BucketWords : kFlow4BucketWords;
    for (int w = 0; w < bw; ++w) rec[w] = 0;
    int t;
    if (!six) {
        rec[0] = f.sip[0];
        rec[1] = f.dip[0];
        rec[2] = f.ports;
        t = 3;
    } else {
        for (int w = 0; w < 4; ++w) {
            rec[w] = f.sip[w];
            rec[4 + w] = f.dip[w];
        }
        rec[8] = f.ports;
        t = 9;
    }
    rec[t] = (f.meta & 0xFFFFFu) | kSlotValid;
    rec[t + 1] = res_hi(f.action, next_nf, f.out_port);
    rec[t + 2] = res_lo(SFC_STAGE_FLOW, f.id);
}

void FlowTable::buckets_of(const uint32_t* rec, uint32_t* b) const {
    uint32_t x, y;
    if (v6) {
        flow6_hash(rec, rec + 4, rec[8], rec[9] & 0xFFFFFu, S, x, y);
        for (int j = 0; j < nhash; ++j) b[j] = flow_bucket(x, y, j, S + 10, nbuckets);
    } else {
        flow4_hash(rec[0], rec[1], rec[2], rec[3] & 0xFFFFFu, S, x, y);
        for (int j = 0; j < nhash; ++j) b[j] = flow_bucket(x, y, j, S + 4, nbuckets);
    }
}

bool FlowTable::insert(const uint32_t* rec, std::vector<uint32_t>* dirty) {
    const int BW = words(), TAG = tag_index();
    uint32_t cur[kFlow6BucketWords], tmp[kFlow6BucketWords];
    memcpy(cur, rec, 4 * BW);
    uint32_t came_from = 0xFFFFFFFFu;
    for (int kick = 0; kick < 4000; ++kick) {
        uint32_t b[kMaxHashes];
        buckets_of(cur, b);
        for (int j = 0; j < nhash; ++j)
            for (int sl = 0; sl < slots; ++sl) {
                const uint32_t idx = b[j] * (uint32_t)slots + (uint32_t)sl;
                uint32_t* p = tab + (size_t)idx * BW;
                if (!(p[TAG] & kSlotValid)) {
                    memcpy(p, cur, 4 * BW);
                    if (dirty) dirty->push_back(idx);
                    return true;
                }
            }
        // Every candidate slot is taken.  Evict, by preference, a resident that is NOT in its first-choice
        // bucket: it goes home (or a step closer), so the walk drifts towards tables in which a lookup ends
        // at the first probe.  Failing that, a random resident as in plain random-walk cuckoo insertion.
        rng = mix64(rng + kGamma);
        uint32_t idx = 0xFFFFFFFFu;
        int j = 0;
        for (int jj = 0; jj < nhash && idx == 0xFFFFFFFFu; ++jj) {
            if (b[jj] == came_from) continue;
            for (int sl = 0; sl < slots; ++sl) {
                const uint32_t cand = b[jj] * (uint32_t)slots + (uint32_t)sl;
                uint32_t hb[kMaxHashes];
                buckets_of(tab + (size_t)cand * BW, hb);
                if (hb[0] != b[jj] && (rng >> (8 + sl)) % 8 != 0) {  // a little randomness keeps walks from cycling
                    idx = cand;
                    j = jj;
                    break;
                }
            }
        }
        if (idx == 0xFFFFFFFFu) {
            j = (int)(rng % (uint64_t)nhash);
            if (b[j] == came_from) j = (j + 1) % nhash;
            idx = b[j] * (uint32_t)slots + (uint32_t)((rng >> 32) % (uint64_t)slots);
        }
        uint32_t* p = tab + (size_t)idx * BW;
        memcpy(tmp, p, 4 * BW);
        memcpy(p, cur, 4 * BW);
        memcpy(cur, tmp, 4 * BW);
        if (dirty) dirty->push_back(idx);
        came_from = b[j];
    }
    return false;  // the caller rebuilds the table (new seed / more buckets)
}

bool FlowTable::erase(const uint32_t* rec, std::vector<uint32_t>* dirty) {
    const int BW = words(), TAG = tag_index();
    uint32_t b[kMaxHashes];
    buckets_of(rec, b);
    for (int j = 0; j < nhash; ++j)
        for (int sl = 0; sl < slots; ++sl) {
            const uint32_t idx = b[j] * (uint32_t)slots + (uint32_t)sl;
            uint32_t* p = tab + (size_t)idx * BW;
            if ((p[TAG] & kSlotValid) && p[TAG] == rec[TAG] && !memcmp(p, rec, 4 * TAG)) {  // key words precede the tag
                memset(p, 0, 4 * BW);
                if (dirty) dirty->push_back(idx);
                return true;
            }
        }
    return false;
}

uint32_t FlowTable::count_first() const {
    const int BW = words(), TAG = tag_index();
    uint32_t n = 0, b[kMaxHashes];
    for (uint64_t i = 0; i < (uint64_t)nbuckets * slots; ++i) {
        const uint32_t* p = tab + (size_t)i * BW;
        if (!(p[TAG] & kSlotValid)) continue;
        buckets_of(p, b);
        n += b[0] == (uint32_t)(i / slots);
    }
    return n;
}


// ------------------------------------------------------------------ IndexedFlows ---------
namespace {
// one or two record units for a flow (format 2); returns the number of units
int encode_units(const FlowRec& f, uint32_t next_nf, uint32_t* u) {
    const bool six = f.ip_ver == 6;
    for (int w = 0; w < 2 * kFrecUnitWords; ++w) u[w] = 0;
    u[0] = (f.meta & 0xFFFFFu) | kSlotValid | (six ? kSlotV6 : 0u);
    u[1] = f.ports;
    u[2] = res_hi(f.action, next_nf, f.out_port);
    u[3] = res_lo(SFC_STAGE_FLOW, f.id);
    if (!six) {
        u[4] = f.sip[0];
        u[5] = f.dip[0];
        return 1;
    }
    for (int w = 0; w < 4; ++w) {
        u[4 + w] = f.sip[w];
        u[8 + w] = f.dip[w];
    }
    return 2;
}
}  // namespace

IndexedFlows::Where IndexedFlows::where(const uint32_t* sip, const uint32_t* dip, uint32_t ports, uint32_t meta20,
                                        bool v6) const {
    uint32_t x, y;
    flow6_hash(sip, dip, ports, flow2_meta(meta20, v6), S, x, y);
    Where w;
    w.b0 = flow_bucket(x, y, 0, S + 10, nbuckets);
    w.b1 = flow_bucket(x, y, 1, S + 10, nbuckets);
    w.fp = flow_fp(x, y, S + 10) >> bits;
    return w;
}

IndexedFlows::Where IndexedFlows::where_of_unit(uint32_t unit) const {
    const uint32_t* u = rec + (size_t)unit * kFrecUnitWords;
    const bool six = (u[0] & kSlotV6) != 0;
    uint32_t sip[4] = {u[4], 0, 0, 0}, dip[4] = {u[5], 0, 0, 0};
    if (six)
        for (int w = 0; w < 4; ++w) {
            sip[w] = u[4 + w];
            dip[w] = u[8 + w];
        }
    return where(sip, dip, u[1], u[0] & 0xFFFFFu, six);
}

int64_t IndexedFlows::find(const FlowRec& f, uint32_t* bucket, int* slot) const {
    if (!nbuckets) return -1;
    const bool six = f.ip_ver == 6;
    const Where w = where(f.sip, f.dip, f.ports, f.meta & 0xFFFFFu, six);
    const uint32_t mask = (1u << bits) - 1u, tag = (f.meta & 0xFFFFFu) | kSlotValid | (six ? kSlotV6 : 0u);
    const uint32_t bs[2] = {w.b0, w.b1};
    for (int j = 0; j < (w.b1 == w.b0 ? 1 : 2); ++j) {
        const uint32_t* b = idx + (size_t)bs[j] * kFidxBucketWords;
        for (int s = 0; s < kFidxEntries; ++s) {
            const uint32_t d = b[s] ^ (w.fp << bits);
            if (d - 1u >= mask) continue;  // empty, or another fingerprint
            const uint32_t unit = d - 1u;
            const uint32_t* u = rec + (size_t)unit * kFrecUnitWords;
            bool eq = u[0] == tag && u[1] == f.ports;
            if (!six) eq = eq && u[4] == f.sip[0] && u[5] == f.dip[0];
            else
                for (int k = 0; k < 4; ++k) eq = eq && u[4 + k] == f.sip[k] && u[8 + k] == f.dip[k];
            if (eq) {
                if (bucket) *bucket = bs[j];
                if (slot) *slot = s;
                return unit;
            }
        }
    }
    return -1;
}

long IndexedFlows::room() const {
    const long by_index = (long)nbuckets * kFidxEntries * 85 / 100 - (long)count;
    const long by_units = ((long)units - (long)used) / 2 + (long)free2.size();  // worst case: every new flow is IPv6
    return std::min(by_index, by_units);
}

// Put the entry for `unit` into one of its two buckets, moving residents out of the way if both are full.
bool IndexedFlows::place(uint32_t unit, Where w, std::vector<uint32_t>* dirty) {
    uint32_t came_from = 0xFFFFFFFFu;
    for (int kick = 0; kick < 500; ++kick) {
        const uint32_t e = (w.fp << bits) | (unit + 1u);
        const uint32_t bs[2] = {w.b0, w.b1};
        for (int j = 0; j < (w.b1 == w.b0 ? 1 : 2); ++j) {
            uint32_t* b = idx + (size_t)bs[j] * kFidxBucketWords;
            for (int s = 0; s < kFidxEntries; ++s)
                if (!b[s]) {
                    b[s] = e;
                    touch_idx(bs[j], dirty);
                    if (j == 1) {
                        ++idx[(size_t)w.b0 * kFidxBucketWords + kFidxEntries];
                        touch_idx(w.b0, dirty);
                        ++second;
                    }
                    return true;
                }
        }
        // both full: take the place of a resident that can move, and go on with that one
        rng = mix64(rng + kGamma);
        uint32_t B = (rng & 1) ? w.b1 : w.b0;
        if (B == came_from) B = B == w.b0 ? w.b1 : w.b0;
        uint32_t* b = idx + (size_t)B * kFidxBucketWords;
        int s = -1;
        Where vw{};
        uint32_t vunit = 0;
        for (int t = 0; t < kFidxEntries && s < 0; ++t) {
            const int cand = (int)(((rng >> 8) + (uint64_t)t) % kFidxEntries);
            vunit = (b[cand] & ((1u << bits) - 1u)) - 1u;
            vw = where_of_unit(vunit);
            if (vw.b0 != vw.b1) s = cand;
        }
        if (s < 0) return false;
        b[s] = e;
        touch_idx(B, dirty);
        if (B != w.b0) {  // the newcomer lives in its second bucket
            ++idx[(size_t)w.b0 * kFidxBucketWords + kFidxEntries];
            touch_idx(w.b0, dirty);
            ++second;
        }
        if (B != vw.b0) {  // the resident leaves its second bucket
            --idx[(size_t)vw.b0 * kFidxBucketWords + kFidxEntries];
            touch_idx(vw.b0, dirty);
            --second;
        }
        // the resident must now go to its other bucket: express that as "both choices are that bucket"
        const uint32_t other = B == vw.b0 ? vw.b1 : vw.b0;
        const bool other_is_second = other != vw.b0;
        came_from = B;
        unit = vunit;
        // try the other bucket directly; if it is full the loop kicks from there
        uint32_t* ob = idx + (size_t)other * kFidxBucketWords;
        bool placed = false;
        for (int t = 0; t < kFidxEntries && !placed; ++t)
            if (!ob[t]) {
                ob[t] = (vw.fp << bits) | (vunit + 1u);
                touch_idx(other, dirty);
                placed = true;
            }
        if (placed) {
            if (other_is_second) {
                ++idx[(size_t)vw.b0 * kFidxBucketWords + kFidxEntries];
                touch_idx(vw.b0, dirty);
                ++second;
            }
            return true;
        }
        w = vw;  // continue: the evicted flow looks for room, not back where it came from
    }
    return false;
}

bool IndexedFlows::insert(const FlowRec& f, uint32_t next_nf, std::vector<uint32_t>* dirty) {
    uint32_t u[2 * kFrecUnitWords];
    const int n = encode_units(f, next_nf, u);
    uint32_t unit;
    if (n == 1 && !free1.empty()) {
        unit = free1.back();
        free1.pop_back();
    } else if (n == 2 && !free2.empty()) {
        unit = free2.back();
        free2.pop_back();
    } else if (n == 1 && !free2.empty()) {  // split a released pair
        unit = free2.back();
        free2.pop_back();
        free1.push_back(unit + 1);
    } else {
        if (used + (uint32_t)n > units) return false;
        unit = used;
        used += (uint32_t)n;
    }
    memcpy(rec + (size_t)unit * kFrecUnitWords, u, (size_t)n * kFrecUnitWords * 4);
    for (int k = 0; k < n; ++k) touch_rec(unit + (uint32_t)k, dirty);
    ++count;
    return place(unit, where(f.sip, f.dip, f.ports, f.meta & 0xFFFFFu, f.ip_ver == 6), dirty);
}

bool IndexedFlows::erase(const FlowRec& f, std::vector<uint32_t>* dirty) {
    uint32_t bucket = 0;
    int slot = 0;
    const int64_t unit = find(f, &bucket, &slot);
    if (unit < 0) return false;
    const Where w = where(f.sip, f.dip, f.ports, f.meta & 0xFFFFFu, f.ip_ver == 6);
    idx[(size_t)bucket * kFidxBucketWords + slot] = 0;
    touch_idx(bucket, dirty);
    if (bucket != w.b0) {
        --idx[(size_t)w.b0 * kFidxBucketWords + kFidxEntries];
        touch_idx(w.b0, dirty);
        --second;
    }
    const int n = f.ip_ver == 6 ? 2 : 1;
    memset(rec + (size_t)unit * kFrecUnitWords, 0, (size_t)n * kFrecUnitWords * 4);
    for (int k = 0; k < n; ++k) touch_rec((uint32_t)unit + (uint32_t)k, dirty);
    (n == 2 ? free2 : free1).push_back((uint32_t)unit);
    --count;
    return true;
}

// ------------------------------------------------------------------ image assembly -------
namespace {

struct Blob {
    std::vector<uint8_t>& img;
    explicit Blob(std::vector<uint8_t>& v) : img(v) {}
    uint64_t reserve(size_t bytes, uint8_t fill) {
        size_t off = (img.size() + kImageAlign - 1) / kImageAlign * kImageAlign;
        img.resize(off, 0);
        img.resize(off + bytes, fill);
        return off;
    }
    template <class T>
    T* at(uint64_t off) {
        return reinterpret_cast<T*>(img.data() + off);
    }
};

// ---- cuckoo ------------------------------------------------------------------------------
bool build_cuckoo(Blob& blob, const RuleStore& rs, bool v6, const std::vector<const FlowRec*>& recs, uint64_t& off,
                  uint32_t& nbuckets, uint32_t& slots_out, uint64_t& seed_out, uint32_t& first_out, std::string& err) {
    const int BW = v6 ? kFlow6BucketWords : kFlow4BucketWords;
    size_t n = recs.size();
    int nhash = rs.flow_hashes;
    int max_load = nhash == 2 ? 45 : 85;
    int load_pct = std::min(std::max(rs.flow_load_pct ? rs.flow_load_pct : 40, 5), max_load);
    // Slots per bucket.  One slot = one sector per probe, which is what counts when the table lives in
    // DRAM (B200 sustains ~50 G random sectors/s there whatever their size: tools/randgather.cu) and
    // costs nothing when it is small; it is the default.  Two slots at 80 % load (option "flow_slots" = 2)
    // halve the footprint and pay off in one narrow case — uniform traffic over 1-2 Mi IPv4 flows, where
    // the smaller table is still largely in L2 (2 Mi flows: 19.6 vs 24.3 us per Mi packets) — but lose
    // on skewed traffic, whose hot buckets sit in L2 either way (configs[2]: 28.8 vs 22.2 us).
    int slots = rs.flow_slots == 2 ? 2 : 1;
    if (slots == 2) load_pct = std::max(load_pct, (int)kFlowDenseLoadPct);
    slots_out = (uint32_t)slots;
    uint64_t nb = std::max<uint64_t>(8, (n * 100 + (uint64_t)load_pct * slots - 1) / ((uint64_t)load_pct * slots));
    std::vector<uint8_t> nf(SFC_MAX_PORTS, (uint8_t)SFC_NF_NONE);
    for (uint32_t p = 0; p < SFC_MAX_PORTS; ++p) nf[p] = nf_of_port(rs, p);
    for (int grow = 0; grow < 6; ++grow, nb = nb + nb / 8 + 1) {
        if (nb > 0xFFFFFFF0ull) break;
        for (int attempt = 0; attempt < 4; ++attempt) {
            size_t start = blob.img.size();
            off = blob.reserve((size_t)nb * slots * BW * 4, 0x00);
            FlowTable c;
            c.v6 = v6;
            c.tab = blob.at<uint32_t>(off);
            c.nbuckets = (uint32_t)nb;
            c.slots = slots;
            c.nhash = nhash;
            uint64_t seed = mix64(0x5FC0C0DEull + 977ull * attempt + 31ull * grow + (v6 ? 7ull : 0ull));
            expand_seed(seed, c.S, v6 ? kFlow6SeedWords : kFlow4SeedWords);
            bool ok = true;
            uint32_t rec[kFlow6BucketWords];
            for (size_t i = 0; i < n && ok; ++i) {
                const FlowRec& f = *recs[i];
                FlowTable::encode(f, f.action == SFC_ACT_FORWARD ? nf[f.out_port] : (uint32_t)SFC_NF_NONE, rec);
                ok = c.insert(rec, nullptr);
            }
            if (ok) {
                nbuckets = (uint32_t)nb;
                seed_out = seed;
                first_out = c.count_first();
                return true;
            }
            blob.img.resize(start);
        }
    }
    err = "flow table: cuckoo insertion did not converge";
    return false;
}

// ---- indexed flow table (format 2) ---------------------------------------------------------
bool build_indexed(Blob& blob, const RuleStore& rs, ImageHeader& H, std::string& err) {
    const size_t n = rs.flows.size();
    uint64_t units_needed = 0;
    for (auto& f : rs.flows) units_needed += f.ip_ver == 6 ? 2 : 1;
    int load_pct = rs.flow_load_pct ? std::min(std::max(rs.flow_load_pct, 5), 85) : 65;
    uint64_t nb = std::max<uint64_t>(1, (n * 100 + (uint64_t)load_pct * kFidxEntries - 1) / ((uint64_t)load_pct * kFidxEntries));
    const uint64_t units = units_needed + units_needed / 8 + 8192;  // room for flows added later
    uint32_t bits = 1;
    while ((1ull << bits) <= units + 1) ++bits;  // entries hold unit + 1
    std::vector<uint8_t> nf(SFC_MAX_PORTS, (uint8_t)SFC_NF_NONE);
    for (uint32_t p = 0; p < SFC_MAX_PORTS; ++p) nf[p] = nf_of_port(rs, p);
    for (int attempt = 0; attempt < 6; ++attempt, nb = nb + nb / 8 + 1) {
        size_t start = blob.img.size();
        H.fidx_off = blob.reserve((size_t)nb * kFidxBucketWords * 4, 0);
        H.frec_off = blob.reserve((size_t)(units + 1) * kFrecUnitWords * 4, 0);
        IndexedFlows t;
        t.idx = blob.at<uint32_t>(H.fidx_off);
        t.rec = blob.at<uint32_t>(H.frec_off);
        t.nbuckets = (uint32_t)nb;
        t.bits = bits;
        t.units = (uint32_t)units;
        const uint64_t seed = mix64(0x5FC0C0DEull + 131ull * attempt + 3ull);
        expand_seed(seed, t.S, kFlow6SeedWords);
        bool ok = true;
        for (size_t i = 0; i < n && ok; ++i) {
            const FlowRec& f = rs.flows[i];
            ok = t.insert(f, f.action == SFC_ACT_FORWARD ? nf[f.out_port] : (uint32_t)SFC_NF_NONE, nullptr);
        }
        if (ok) {
            H.fidx_nbuckets = (uint32_t)nb;
            H.fidx_bits = bits;
            H.fidx_seed = seed;
            H.frec_units = (uint32_t)units;
            H.frec_used = t.used;
            H.fidx_first = t.count - t.second;
            return true;
        }
        blob.img.resize(start);
    }
    err = "flow table: index insertion did not converge";
    return false;
}

// ---- LPM ---------------------------------------------------------------------------------
inline uint32_t addr_bits(const uint32_t* w, int pos, int n) {  // n <= 16, MSB-first bit string
    uint64_t two = ((uint64_t)w[pos >> 5] << 32) | ((pos >> 5) < 3 ? w[(pos >> 5) + 1] : 0u);
    int sh = 64 - (pos & 31) - n;
    return (uint32_t)(two >> sh) & ((1u << n) - 1u);
}

// 16-8-8-... multibit trie with leaf pushing, built top-down from the prefix set so that a sub-trie
// holding one single prefix becomes an LpmTail record instead of a chain of 256-entry nodes (random
// IPv6 /48-/64 prefixes would otherwise cost six dependent loads and 6 KB each).
struct Trie {
    struct P {
        const uint32_t* w;
        int len;
        uint32_t val;
    };
    std::vector<uint32_t> root = std::vector<uint32_t>(65536, 0);
    std::vector<uint32_t> nodes;  // 256 entries per node
    std::vector<LpmTail> tails;

    // the entry for a slot whose path is decided through `pos` bits; every prefix of S is longer than pos
    // and lies on that path; F is what the slot inherits from shorter prefixes.  S is sorted by length.
    uint32_t subtree(const std::vector<P>& S, int pos, uint32_t F) {
        if (S.empty()) return F;
        if (S.size() == 1) {
            LpmTail t;
            memset(&t, 0, sizeof(t));
            for (int i = 0; i < 4; ++i) {
                int bits = S[0].len - 32 * i;
                t.w[i] = bits <= 0 ? 0u : (bits >= 32 ? S[0].w[i] : (S[0].w[i] & ~((1u << (32 - bits)) - 1u)));
            }
            t.len = (uint32_t)S[0].len;
            t.match = S[0].val;
            t.fallback = F;
            tails.push_back(t);
            return kLpmTail | (uint32_t)(tails.size() - 1);
        }
        const size_t me = nodes.size() / 256;
        nodes.resize(nodes.size() + 256, F);
        std::map<uint32_t, std::vector<P>> deeper;
        for (const P& p : S) {  // ascending length: longer prefixes overwrite shorter ones
            if (p.len <= pos + 8) {
                int free_bits = pos + 8 - p.len;
                uint32_t first = addr_bits(p.w, pos, 8) & ~((1u << free_bits) - 1u);
                for (uint32_t i = 0; i < (1u << free_bits); ++i) nodes[me * 256 + first + i] = p.val;
            } else {
                deeper[addr_bits(p.w, pos, 8)].push_back(p);
            }
        }
        for (auto& kv : deeper) {
            uint32_t e = subtree(kv.second, pos + 8, nodes[me * 256 + kv.first]);
            nodes[me * 256 + kv.first] = e;
        }
        return kLpmChild | (uint32_t)me;
    }
    void build(const std::vector<P>& all) {  // sorted by length
        std::map<uint32_t, std::vector<P>> deeper;
        for (const P& p : all) {
            if (p.len <= 16) {
                uint32_t first = p.len ? (addr_bits(p.w, 0, 16) & ~((1u << (16 - p.len)) - 1u)) : 0u;
                for (uint32_t i = 0; i < (1u << (16 - p.len)); ++i) root[first + i] = p.val;
            } else {
                deeper[addr_bits(p.w, 0, 16)].push_back(p);
            }
        }
        for (auto& kv : deeper) root[kv.first] = subtree(kv.second, 16, root[kv.first]);
    }
};

void build_lpm(Blob& blob, const RuleStore& rs, int af, const std::map<RuleStore::LpmKey, LpmRec>& recs_map, uint64_t& root_off, uint64_t& nodes_off,
               uint64_t& nh_off, uint32_t& nnodes, uint32_t& count, uint64_t& tail_off, uint32_t& ntails) {
    std::vector<const LpmRec*> sorted;
    std::vector<LpmRec> recs;
    recs.reserve(recs_map.size());
    for (auto& kv : recs_map) recs.push_back(kv.second);
    sorted.reserve(recs.size());
    for (auto& r : recs) sorted.push_back(&r);
    std::stable_sort(sorted.begin(), sorted.end(),
                     [](const LpmRec* a, const LpmRec* b) { return a->len < b->len; });
    Trie t;
    uint32_t max_id = 0;
    std::vector<Trie::P> all;
    all.reserve(sorted.size());
    for (auto* r : sorted) {
        all.push_back({r->w, r->len, r->id + 1});
        max_id = std::max(max_id, r->id);
    }
    t.build(all);
    count = (uint32_t)recs.size();
    nnodes = (uint32_t)(t.nodes.size() / 256);
    root_off = blob.reserve(65536 * 4, 0);
    memcpy(blob.at<uint8_t>(root_off), t.root.data(), 65536 * 4);
    nodes_off = blob.reserve(std::max<size_t>(t.nodes.size() * 4, 4), 0);
    if (!t.nodes.empty()) memcpy(blob.at<uint8_t>(nodes_off), t.nodes.data(), t.nodes.size() * 4);
    ntails = (uint32_t)t.tails.size();
    tail_off = blob.reserve(std::max<size_t>(t.tails.size() * sizeof(LpmTail), 32), 0);
    if (!t.tails.empty()) memcpy(blob.at<uint8_t>(tail_off), t.tails.data(), t.tails.size() * sizeof(LpmTail));
    size_t nh_n = recs.empty() ? 1 : (size_t)max_id + 1;
    nh_off = blob.reserve(nh_n * sizeof(LpmRes), 0);
    for (auto& r : recs) {
        LpmRes& e = blob.at<LpmRes>(nh_off)[r.id];
        e.res_hi = res_hi(SFC_ACT_FORWARD, nf_of_port(rs, r.out_port), r.out_port);
        e.res_lo = res_lo(af == 4 ? SFC_STAGE_LPM4 : SFC_STAGE_LPM6, r.id);
    }
    for (uint32_t i = 0; i < ntails; ++i) {  // a tail hit carries its own result (res_lo = stage | id is computed)
        LpmTail& t = blob.at<LpmTail>(tail_off)[i];
        t.match_hi = blob.at<LpmRes>(nh_off)[t.match - 1].res_hi;
    }
}

// ---- ACL ---------------------------------------------------------------------------------
// Bit-selecting decision trees (in the family of HiCuts / EffiCuts / ByteCuts, cutting on single
// key bits so that ARBITRARY ternary masks stay exact: a rule that does not care about a chosen bit
// goes to both children).  Rules are ranked by (priority desc, id asc); a leaf lists its candidate
// rules in rank order and the kernel takes the first that matches in full, so the trees only ever
// narrow the candidate set — a wrong cut can cost time, never correctness.
//  * a node tests up to 6 bits (64 children; 5 or 4 when memory is short), chosen greedily to minimise
//    the largest child;
//  * below a rule that matches the whole subspace of a node nothing of lower rank can win: pruned;
//  * identical sub-problems (same rule list) are built once and shared;
//  * rules are first split into categories by which fields they leave (almost) open — destination,
//    source, destination port — one tree per populated category, because a rule that is wild in the
//    field a tree cuts on would be copied into every branch (the EffiCuts observation).
void acl_key_words(const sfc_acl_rule& r, uint32_t* val, uint32_t* mask) {
    uint8_t v[kAclKeyBytes], m[kAclKeyBytes];
    memset(v, 0, sizeof(v));
    memset(m, 0, sizeof(m));
    auto put16 = [&](int i, uint16_t x, uint16_t mm) {
        v[i] = x >> 8;
        v[i + 1] = x & 0xFF;
        m[i] = mm >> 8;
        m[i + 1] = mm & 0xFF;
    };
    put16(0, r.in_port, r.in_port_mask);
    put16(2, r.vlan, r.vlan_mask);
    v[4] = r.ip_ver;
    m[4] = r.ip_ver ? 0xFF : 0;
    v[5] = r.proto;
    m[5] = r.proto_mask;
    put16(6, r.sport, r.sport_mask);
    put16(8, r.dport, r.dport_mask);
    for (int i = 0; i < 16; ++i) {
        v[12 + i] = r.src[i];
        m[12 + i] = r.src_mask[i];
        v[28 + i] = r.dst[i];
        m[28 + i] = r.dst_mask[i];
    }
    for (int w = 0; w < kAclKeyWords; ++w) {
        mask[w] = be32(m + 4 * w);
        val[w] = be32(v + 4 * w) & mask[w];
    }
}

struct AclBuilder {
    static constexpr int kMaxDepth = 12;
    int kLeafMax = 8;  // a sub-problem this small becomes a leaf
    struct R {
        uint32_t val[kAclKeyWords], mask[kAclKeyWords];
    };
    int bits_per_node = 5;
    bool one_word = false;
    std::vector<R> rules;           // by rank
    std::vector<AclEntry> nodes;    // child arrays, 2^bits entries each
    std::vector<uint16_t> leaves;   // overflow lists: count, candidates beyond the sixth
    std::map<std::vector<uint16_t>, AclEntry> node_memo, leaf_memo;
    size_t budget_bytes = ~(size_t)0;
    bool over_budget = false;

    static bool bit(const uint32_t* w, int b) { return (w[b >> 5] >> (31 - (b & 31))) & 1u; }
    size_t bytes() const { return nodes.size() * sizeof(AclEntry) + leaves.size() * 2; }

    uint16_t candidate(uint16_t r) const {
        const uint32_t* m = rules[r].mask;
        uint16_t e = r;
        if (m[0] | m[8] | m[9] | m[10]) e |= kAclNeedG1;
        if (m[4] | m[5] | m[6]) e |= kAclNeedG2;
        return e;
    }
    AclEntry make_leaf(const std::vector<uint16_t>& rs) {
        auto it = leaf_memo.find(rs);
        if (it != leaf_memo.end()) return it->second;
        uint16_t c[kAclInline];
        for (int i = 0; i < kAclInline; ++i) c[i] = i < (int)rs.size() ? candidate(rs[i]) : kAclNone;
        AclEntry e;
        e.x = kAclLeafBit;
        if ((int)rs.size() > kAclInline) {
            e.x |= kAclMoreBit | (uint32_t)leaves.size();
            leaves.push_back((uint16_t)(rs.size() - kAclInline));
            for (size_t i = kAclInline; i < rs.size(); ++i) leaves.push_back(candidate(rs[i]));
        }
        e.y = c[0] | (uint32_t)c[1] << 16;
        e.z = c[2] | (uint32_t)c[3] << 16;
        e.w = c[4] | (uint32_t)c[5] << 16;
        return leaf_memo[rs] = e;
    }

    AclEntry build(std::vector<uint16_t> rs, const uint32_t* fixed, int depth) {
        // prune everything ranked below the first rule that covers this node's whole subspace
        for (size_t i = 0; i < rs.size(); ++i) {
            bool covered = true;
            for (int w = 0; w < kAclKeyWords && covered; ++w) covered = (rules[rs[i]].mask[w] & ~fixed[w]) == 0;
            if (covered) {
                rs.resize(i + 1);
                break;
            }
        }
        if ((int)rs.size() <= kLeafMax || depth >= kMaxDepth || over_budget) return make_leaf(rs);
        // Sub-problems are shared when the rule list AND the already-decided bits that any of these
        // rules cares about are the same: the "covers the whole subspace" pruning further down
        // depends on exactly that, so a subtree must not be reused under different decided bits.
        std::vector<uint16_t> memo_key = rs;
        {
            uint32_t care[kAclKeyWords] = {0};
            for (uint16_t r : rs)
                for (int w = 0; w < kAclKeyWords; ++w) care[w] |= rules[r].mask[w];
            for (int w = 0; w < kAclKeyWords; ++w) {
                uint32_t f = fixed[w] & care[w];
                memo_key.push_back((uint16_t)f);
                memo_key.push_back((uint16_t)(f >> 16));
            }
        }
        auto memo = node_memo.find(memo_key);
        if (memo != node_memo.end()) return memo->second;

        std::vector<std::vector<uint16_t>> cells(1, rs);
        int chosen[kAclMaxNodeBits], nchosen = 0;
        std::vector<uint32_t> n0(kAclKeyBits), n1(kAclKeyBits), worst(kAclKeyBits);
        std::vector<uint8_t> progress(kAclKeyBits);
        for (int j = 0; j < bits_per_node; ++j) {
            std::fill(worst.begin(), worst.end(), 0u);
            std::fill(progress.begin(), progress.end(), 0);
            for (auto& c : cells) {
                if (c.empty()) continue;
                std::fill(n0.begin(), n0.end(), 0u);
                std::fill(n1.begin(), n1.end(), 0u);
                for (uint16_t r : c)
                    for (int w = 0; w < kAclKeyWords; ++w)
                        for (uint32_t m = rules[r].mask[w]; m; m &= m - 1) {
                            int bpos = w * 32 + __builtin_clz(m & (~m + 1));  // lowest set bit of m, as an MSB-first position
                            ((rules[r].val[w] >> (31 - (bpos & 31))) & 1u ? n1 : n0)[bpos]++;
                        }
                for (int b = 0; b < kAclKeyBits; ++b) {
                    uint32_t nw = (uint32_t)c.size() - n0[b] - n1[b];
                    uint32_t cost = std::max(n0[b], n1[b]) + nw;  // size of the larger child
                    if (cost < c.size()) progress[b] = 1;
                    worst[b] = std::max(worst[b], cost);
                }
            }
            int best = -1;
            for (int b = 0; b < kAclKeyBits; ++b) {
                if (!progress[b] || bit(fixed, b)) continue;
                if (one_word && nchosen && (b >> 5) != (chosen[0] >> 5)) continue;  // all bits of a node from one key word
                bool used = false;
                for (int q = 0; q < nchosen; ++q) used = used || chosen[q] == b;
                if (!used && (best < 0 || worst[b] < worst[best])) best = b;
            }
            if (best < 0) break;
            chosen[nchosen++] = best;
            std::vector<std::vector<uint16_t>> next;
            next.reserve(cells.size() * 2);
            for (auto& c : cells) {
                std::vector<uint16_t> c0, c1;
                for (uint16_t r : c) {
                    bool cares = bit(rules[r].mask, best), one = bit(rules[r].val, best);
                    if (!cares || !one) c0.push_back(r);
                    if (!cares || one) c1.push_back(r);
                }
                next.push_back(std::move(c0));
                next.push_back(std::move(c1));
            }
            cells.swap(next);
        }
        if (!nchosen) return make_leaf(rs);

        const size_t base = nodes.size();
        nodes.resize(base + ((size_t)1 << nchosen), make_leaf({}));
        if (bytes() > budget_bytes) over_budget = true;
        // fields: unused ones (reading the zero row) first, then the chosen bits, first chosen = most significant
        uint16_t field[kAclMaxNodeBits];
        for (int q = 0; q < kAclMaxNodeBits; ++q) {
            int c = q - (kAclMaxNodeBits - nchosen);
            field[q] = c < 0 ? (uint16_t)kAclZeroField : (uint16_t)((chosen[c] >> 5) * 128 + (31 - (chosen[c] & 31)));
        }
        AclEntry entry;
        entry.x = (uint32_t)base;
        entry.y = field[0] | (uint32_t)field[1] << 16;
        entry.z = field[2] | (uint32_t)field[3] << 16;
        entry.w = field[4] | (uint32_t)field[5] << 16;
        node_memo[memo_key] = entry;
        uint32_t f[kAclKeyWords];
        memcpy(f, fixed, sizeof(f));
        for (int q = 0; q < nchosen; ++q) f[chosen[q] >> 5] |= 1u << (31 - (chosen[q] & 31));
        for (int idx = 0; idx < (1 << nchosen); ++idx) {
            AclEntry child = build(cells[idx], f, depth + 1);
            nodes[base + idx] = child;
        }
        return entry;
    }
};

void build_acl(Blob& blob, const std::vector<AclRec>& recs, ImageHeader& H) {
    std::vector<const AclRec*> sorted;
    for (auto& r : recs) sorted.push_back(&r);
    std::stable_sort(sorted.begin(), sorted.end(), [](const AclRec* a, const AclRec* b) {
        if (a->r.priority != b->r.priority) return a->r.priority > b->r.priority;
        return a->id < b->id;
    });
    uint32_t n = (uint32_t)sorted.size();
    H.acl_nrules = n;
    H.acl_ntrees = 0;
    if (!n) {
        H.acl_nodes_off = H.acl_leaf_off = H.acl_rule_off = H.acl_res_off = blob.reserve(32, 0);
        H.acl_nentries = H.acl_nleafwords = 0;
        return;
    }
    std::vector<AclBuilder::R> rules(n);
    for (uint32_t i = 0; i < n; ++i) acl_key_words(sorted[i]->r, rules[i].val, rules[i].mask);
    auto popcnt = [](const uint32_t* w, int from, int to) {
        int c = 0;
        for (int i = from; i < to; ++i) c += __builtin_popcount(w[i]);
        return c;
    };
    // Try the cheapest lookups first: one tree over all rules, then forests split by which of
    // (dst, dport, src) a rule leaves open; take the first that fits the memory budget.
    // Wider nodes mean fewer dependent loads; they are tried first and given up for memory.
    const int feature_sets[6] = {0, 0, 0, 1, 5, 7}, node_bits[6] = {6, 5, 4, 4, 4, 4};
    const char* force = getenv("SFC_ACL_FEAT");
    const char* force_bits = getenv("SFC_ACL_BITS");
    const char* budget_env = getenv("SFC_ACL_BUDGET_MB");
    const size_t budget = (budget_env ? (size_t)atoi(budget_env) : 40) << 20;
    AclBuilder B;
    for (int attempt = 0; attempt < 6; ++attempt) {
        const int feat = force ? atoi(force) : feature_sets[attempt];
        B = AclBuilder();
        B.rules = rules;
        B.bits_per_node = force_bits ? std::min(std::max(atoi(force_bits), 1), kAclMaxNodeBits) : (force ? 4 : node_bits[attempt]);
        B.budget_bytes = (force || attempt == 5) ? ~(size_t)0 : budget;
        B.one_word = getenv("SFC_ACL_ONEWORD") != nullptr;
        if (const char* lf = getenv("SFC_ACL_LEAF")) B.kLeafMax = std::min(std::max(atoi(lf), 1), 64);
        std::vector<std::vector<uint16_t>> cat(8);
        for (uint32_t i = 0; i < n; ++i) {
            const uint32_t* m = rules[i].mask;
            int c = ((popcnt(m, 7, 11) < 8 ? 1 : 0) | (popcnt(m, 3, 7) < 8 ? 2 : 0) | (__builtin_popcount(m[2]) < 16 ? 4 : 0)) & feat;
            cat[c].push_back((uint16_t)i);
        }
        std::vector<uint16_t> misc;  // small categories share one tree
        std::vector<std::vector<uint16_t>> groups;
        for (auto& c : cat) {
            if (c.empty()) continue;
            if (c.size() < 24) misc.insert(misc.end(), c.begin(), c.end());
            else groups.push_back(c);
        }
        if (!misc.empty()) {
            std::sort(misc.begin(), misc.end());
            groups.push_back(misc);
        }
        uint32_t fixed0[kAclKeyWords] = {0};
        H.acl_ntrees = 0;
        for (auto& g : groups) H.acl_roots[H.acl_ntrees++] = B.build(g, fixed0, 0);
        if (!B.over_budget || force) break;
    }
    H.acl_nentries = (uint32_t)B.nodes.size();
    H.acl_nleafwords = (uint32_t)B.leaves.size();
    H.acl_nodes_off = blob.reserve(std::max<size_t>(B.nodes.size() * sizeof(AclEntry), 16), 0);
    if (!B.nodes.empty()) memcpy(blob.at<uint8_t>(H.acl_nodes_off), B.nodes.data(), B.nodes.size() * sizeof(AclEntry));
    H.acl_leaf_off = blob.reserve(std::max<size_t>(B.leaves.size() * 2, 16), 0);
    memcpy(blob.at<uint8_t>(H.acl_leaf_off), B.leaves.data(), B.leaves.size() * 2);
    H.acl_rule_off = blob.reserve((size_t)n * kAclRuleWords * 4, 0);
    H.acl_res_off = blob.reserve((size_t)n * 4, 0);
    uint32_t* rw = blob.at<uint32_t>(H.acl_rule_off);
    uint32_t* res = blob.at<uint32_t>(H.acl_res_off);
    const int g0[4] = {1, 2, 7, 3}, g1[4] = {0, 8, 9, 10}, g2[3] = {4, 5, 6};
    for (uint32_t i = 0; i < n; ++i) {
        uint32_t* o = rw + (size_t)i * kAclRuleWords;
        for (int j = 0; j < 4; ++j) {
            o[j] = rules[i].val[g0[j]];
            o[4 + j] = rules[i].mask[g0[j]];
            o[8 + j] = rules[i].val[g1[j]];
            o[12 + j] = rules[i].mask[g1[j]];
        }
        for (int j = 0; j < 3; ++j) {
            o[16 + j] = rules[i].val[g2[j]];
            o[20 + j] = rules[i].mask[g2[j]];
        }
        res[i] = sorted[i]->id | (sorted[i]->r.verdict == SFC_ACL_DENY ? 0x80000000u : 0u);
    }
}

}  // namespace

// Resolve the OpenFlow priorities per match (exact and wildcard apart) and write the two tables of copy
// `copy`; the header inside `image` gets the new counts and its active-copy fields.  Returns false if
// the exact rules no longer fit (load above one half).
static bool fill_l2_copy(const RuleStore& rs, uint8_t* image, int copy, std::vector<uint32_t>* dirty) {
    ImageHeader H;
    memcpy(&H, image, sizeof(H));
    std::map<std::pair<uint32_t, uint32_t>, const L2Rule*> exact;  // (key_hi,key_lo) -> winner
    std::map<uint16_t, const L2Rule*> wild;
    auto better = [](const L2Rule* a, const L2Rule* b) {  // a beats b
        return a->priority > b->priority || (a->priority == b->priority && a->id < b->id);
    };
    for (auto& r : rs.l2) {
        if (r.has_mac) {
            uint32_t hi = ((uint32_t)r.in_port << 16) | ((uint32_t)r.mac[0] << 8) | r.mac[1];
            uint32_t lo = be32(r.mac + 2);
            auto key = std::make_pair(hi, lo);
            auto it = exact.find(key);
            if (it == exact.end() || better(&r, it->second)) exact[key] = &r;
        } else {
            auto it = wild.find(r.in_port);
            if (it == wild.end() || better(&r, it->second)) wild[r.in_port] = &r;
        }
    }
    if (exact.size() * 2 > H.l2x_cap) return false;
    std::vector<L2Entry> x(H.l2x_cap, L2Entry{0, 0, 0, 0});
    std::vector<L2Wild> w(H.l2w_n, L2Wild{0, 0});
    for (auto& kv : exact) {
        uint32_t i = l2_hash(kv.first.first, kv.first.second) & H.l2x_mask;
        while (x[i].res_hi) i = (i + 1) & H.l2x_mask;
        x[i].key_hi = kv.first.first;
        x[i].key_lo = kv.first.second;
        x[i].res_lo = res_lo(SFC_STAGE_L2_EXACT, kv.second->id);
        x[i].res_hi = res_hi(SFC_ACT_FORWARD, nf_of_port(rs, kv.second->out_port), kv.second->out_port);
    }
    for (auto& kv : wild) {
        if (kv.first >= H.l2w_n) continue;
        w[kv.first].res_hi = res_hi(SFC_ACT_FORWARD, nf_of_port(rs, kv.second->out_port), kv.second->out_port);
        w[kv.first].res_lo = res_lo(SFC_STAGE_L2_WILD, kv.second->id);
    }
    // write sector by sector, reporting the sectors whose bytes changed
    auto put = [&](uint64_t off, const void* src, size_t bytes) {
        const uint8_t* p = (const uint8_t*)src;
        for (size_t o = 0; o < bytes; o += 32) {
            const size_t m = std::min<size_t>(32, bytes - o);
            if (memcmp(image + off + o, p + o, m)) {
                memcpy(image + off + o, p + o, m);
                if (dirty) dirty->push_back((uint32_t)((off + o) / 32));
            }
        }
    };
    put(H.l2x_copy_off[copy], x.data(), x.size() * sizeof(L2Entry));
    put(H.l2w_copy_off[copy], w.data(), w.size() * sizeof(L2Wild));
    H.l2_active = (uint32_t)copy;
    H.l2x_off = H.l2x_copy_off[copy];
    H.l2w_off = H.l2w_copy_off[copy];
    H.l2x_count = (uint32_t)exact.size();
    H.l2w_count = (uint32_t)wild.size();
    memcpy(image, &H, sizeof(H));
    return true;
}

bool rebuild_l2_copy(const RuleStore& rs, uint8_t* image, int copy, std::vector<uint32_t>* dirty) {
    return fill_l2_copy(rs, image, copy, dirty);
}

bool build_image(const RuleStore& rs, uint64_t epoch, std::vector<uint8_t>& image, std::string& err) {
    image.clear();
    image.resize(kImageHeaderBytes, 0);
    Blob blob(image);
    ImageHeader H;
    memset(&H, 0, sizeof(H));
    H.magic = kImageMagic;
    H.version = kImageVersion;
    H.header_bytes = (uint32_t)kImageHeaderBytes;
    H.epoch = epoch;

    // flows, split by version
    std::vector<const FlowRec*> f4, f6;
    for (auto& f : rs.flows) (f.ip_ver == 4 ? f4 : f6).push_back(&f);
    H.flow4_count = (uint32_t)f4.size();
    H.flow6_count = (uint32_t)f6.size();
    H.flow4_nhash = H.flow6_nhash = (uint32_t)rs.flow_hashes;
    // Format: the indexed table keeps what a lookup touches small (7 index bytes and one 32-byte record per
    // IPv4 flow) and serves both families with one code path; the sector tables cost one sector per probe
    // and nothing else, which is what counts once the records cannot stay in L2 anyway.
    const uint64_t units_needed = (uint64_t)f4.size() + 2ull * f6.size();
    uint32_t format = (uint32_t)rs.flow_format;
    // measured on B200 (profiles/r02_experiments.md): the sector tables are faster at every size, so they are
    // what "by size" picks; the indexed table is for images that must be small
    if (format != kFlowFormatSector && format != kFlowFormatIndexed) format = kFlowFormatSector;
    if (format == kFlowFormatIndexed && units_needed + units_needed / 8 + 8192 >= (1ull << 27)) format = kFlowFormatSector;
    H.flow_format = format;
    if (format == kFlowFormatIndexed) {
        std::vector<const FlowRec*> none;
        H.flow4_count = H.flow6_count = 0;  // the sector tables stay empty (8 buckets each); counts are restored below
        if (!build_cuckoo(blob, rs, false, none, H.flow4_off, H.flow4_nbuckets, H.flow4_slots, H.flow4_seed, H.flow4_first, err)) return false;
        if (!build_cuckoo(blob, rs, true, none, H.flow6_off, H.flow6_nbuckets, H.flow6_slots, H.flow6_seed, H.flow6_first, err)) return false;
        H.flow4_count = (uint32_t)f4.size();
        H.flow6_count = (uint32_t)f6.size();
        if (!build_indexed(blob, rs, H, err)) return false;
    } else {
        if (!build_cuckoo(blob, rs, false, f4, H.flow4_off, H.flow4_nbuckets, H.flow4_slots, H.flow4_seed, H.flow4_first, err)) return false;
        if (!build_cuckoo(blob, rs, true, f6, H.flow6_off, H.flow6_nbuckets, H.flow6_slots, H.flow6_seed, H.flow6_first, err)) return false;
        H.fidx_off = H.frec_off = blob.reserve(64, 0);
        H.fidx_nbuckets = 1;
        H.fidx_bits = 1;
        H.frec_units = 1;
    }

    // L2 steering: two copies of the exact and the wildcard table, with room to grow, so that later rule
    // edits are rebuilt into the idle copy instead of rebuilding the image (rebuild_l2_copy)
    {
        size_t n_exact = 0;
        for (auto& r : rs.l2) n_exact += r.has_mac;
        uint32_t cap = 256;
        while (cap < n_exact * 4) cap <<= 1;
        H.l2x_cap = cap;
        H.l2x_mask = cap - 1;
        H.l2w_n = SFC_MAX_PORTS;
        for (int c = 0; c < 2; ++c) {
            H.l2x_copy_off[c] = blob.reserve((size_t)cap * sizeof(L2Entry), 0);
            H.l2w_copy_off[c] = blob.reserve((size_t)H.l2w_n * sizeof(L2Wild), 0);
        }
        memcpy(image.data(), &H, sizeof(H));  // fill_l2_copy reads the capacities from the header in the image
        fill_l2_copy(rs, image.data(), 0, nullptr);
        memcpy(&H, image.data(), sizeof(H));
    }
    build_lpm(blob, rs, 4, rs.lpm4, H.lpm4_root_off, H.lpm4_nodes_off, H.lpm4_nh_off, H.lpm4_nnodes, H.lpm4_count,
              H.lpm4_tail_off, H.lpm4_ntails);
    build_lpm(blob, rs, 6, rs.lpm6, H.lpm6_root_off, H.lpm6_nodes_off, H.lpm6_nh_off, H.lpm6_nnodes, H.lpm6_count,
              H.lpm6_tail_off, H.lpm6_ntails);
    build_acl(blob, rs.acl, H);

    size_t total = (image.size() + kImageAlign - 1) / kImageAlign * kImageAlign;
    image.resize(total, 0);
    H.total_bytes = total;
    memcpy(image.data(), &H, sizeof(H));
    return true;
}

const char* validate_image(const uint8_t* img, size_t len) {
    if (len < kImageHeaderBytes) return "image: shorter than its header";
    ImageHeader H;
    memcpy(&H, img, sizeof(H));
    if (H.magic != kImageMagic || H.version != kImageVersion) return "image: bad magic or version";
    if (H.total_bytes != len || H.header_bytes != kImageHeaderBytes)
        return "image: size field does not match the buffer length";
    auto inside = [&](uint64_t off, uint64_t bytes) { return off >= kImageHeaderBytes && off + bytes <= len; };
    bool ok = (H.flow4_slots == 1 || H.flow4_slots == 2) && (H.flow6_slots == 1 || H.flow6_slots == 2) &&
              inside(H.flow4_off, (uint64_t)H.flow4_nbuckets * H.flow4_slots * kFlow4BucketBytes) &&
              inside(H.flow6_off, (uint64_t)H.flow6_nbuckets * H.flow6_slots * kFlow6BucketBytes) &&
              inside(H.l2x_off, ((uint64_t)H.l2x_mask + 1) * sizeof(L2Entry)) &&
              ((H.l2x_mask & (H.l2x_mask + 1)) == 0) && inside(H.l2w_off, (uint64_t)H.l2w_n * sizeof(L2Wild)) &&
              inside(H.lpm4_root_off, 65536 * 4) && inside(H.lpm4_nodes_off, (uint64_t)H.lpm4_nnodes * 1024) &&
              inside(H.lpm6_root_off, 65536 * 4) && inside(H.lpm6_nodes_off, (uint64_t)H.lpm6_nnodes * 1024) &&
              inside(H.lpm4_nh_off, 8) && inside(H.lpm6_nh_off, 8) &&
              inside(H.lpm4_tail_off, std::max<uint64_t>((uint64_t)H.lpm4_ntails * sizeof(LpmTail), 32)) &&
              inside(H.lpm6_tail_off, std::max<uint64_t>((uint64_t)H.lpm6_ntails * sizeof(LpmTail), 32)) && H.flow4_nhash >= 1 &&
              H.flow4_nhash <= kMaxHashes && H.flow6_nhash >= 1 && H.flow6_nhash <= kMaxHashes && H.flow4_nbuckets &&
              H.flow6_nbuckets && H.acl_nrules <= SFC_MAX_ACL_RULES && H.acl_ntrees <= kAclMaxTrees;
    if (ok && H.acl_nrules)
        ok = inside(H.acl_nodes_off, (uint64_t)H.acl_nentries * sizeof(AclEntry)) &&
             inside(H.acl_leaf_off, (uint64_t)H.acl_nleafwords * 2) &&
             inside(H.acl_rule_off, (uint64_t)H.acl_nrules * kAclRuleWords * 4) &&
             inside(H.acl_res_off, (uint64_t)H.acl_nrules * 4);
    if (!ok) return "image: a section lies outside the buffer or a field is out of range";
    // the kernel reads these sections with 16- and 32-byte loads
    const uint64_t offs32[] = {H.flow4_off, H.flow6_off, H.fidx_off, H.frec_off, H.lpm4_tail_off, H.lpm6_tail_off, H.acl_rule_off};
    const uint64_t offs16[] = {H.l2x_off, H.l2w_off, H.acl_nodes_off, H.lpm4_root_off, H.lpm4_nodes_off, H.lpm6_root_off,
                               H.lpm6_nodes_off, H.lpm4_nh_off, H.lpm6_nh_off, H.acl_leaf_off, H.acl_res_off};
    for (uint64_t o : offs32)
        if (o & 31) return "image: a section is not 32-byte aligned";
    for (uint64_t o : offs16)
        if (o & 15) return "image: a section is not 16-byte aligned";
    // L2 exact table: linear probing ends at an empty entry, so there must be one
    {
        if ((uint64_t)H.l2x_count * 2 > (uint64_t)H.l2x_mask + 1) return "image: the L2 exact table is more than half full";
        const L2Entry* x = (const L2Entry*)(img + H.l2x_off);
        uint32_t used = 0;
        for (uint64_t i = 0; i <= H.l2x_mask; ++i) used += x[i].res_hi != 0;
        if (used > ((uint64_t)H.l2x_mask + 1) / 2) return "image: the L2 exact table has too few empty entries";
    }
    // result words stored in the tables: the counting code indexes small arrays with their action and stage
    auto res_ok = [](uint32_t hi, uint32_t lo) { return (hi >> 24) <= SFC_ACT_DROP && (lo >> 28) <= SFC_STAGE_LPM6; };
    {
        const L2Entry* x = (const L2Entry*)(img + H.l2x_off);
        for (uint64_t i = 0; i <= H.l2x_mask; ++i)
            if (x[i].res_hi && !res_ok(x[i].res_hi, x[i].res_lo)) return "image: an L2 entry holds an unknown action or stage";
        const L2Wild* w = (const L2Wild*)(img + H.l2w_off);
        for (uint32_t i = 0; i < H.l2w_n; ++i)
            if (w[i].res_hi && !res_ok(w[i].res_hi, w[i].res_lo)) return "image: an L2 entry holds an unknown action or stage";
        if (H.flow_format == kFlowFormatSector) {
            const uint32_t* f4 = (const uint32_t*)(img + H.flow4_off);
            for (uint64_t i = 0; i < (uint64_t)H.flow4_nbuckets * H.flow4_slots; ++i)
                if ((f4[i * 8 + 3] & kSlotValid) && !res_ok(f4[i * 8 + 4], f4[i * 8 + 5])) return "image: a flow slot holds an unknown action or stage";
            const uint32_t* f6 = (const uint32_t*)(img + H.flow6_off);
            for (uint64_t i = 0; i < (uint64_t)H.flow6_nbuckets * H.flow6_slots; ++i)
                if ((f6[i * 16 + 9] & kSlotValid) && !res_ok(f6[i * 16 + 10], f6[i * 16 + 11])) return "image: a flow slot holds an unknown action or stage";
        }
    }
    if (H.flow_format != kFlowFormatSector && H.flow_format != kFlowFormatIndexed) return "image: unknown flow table format";
    if (H.flow_format == kFlowFormatIndexed) {
        if (H.fidx_bits < 1 || H.fidx_bits > 27 || !H.fidx_nbuckets || ((uint64_t)H.frec_units + 1) >= (1ull << H.fidx_bits) ||
            !inside(H.fidx_off, (uint64_t)H.fidx_nbuckets * kFidxBucketWords * 4) ||
            !inside(H.frec_off, ((uint64_t)H.frec_units + 1) * kFrecUnitWords * 4))
            return "image: the flow index or record section is out of range";
        const uint32_t* e = (const uint32_t*)(img + H.fidx_off);
        const uint32_t mask = (1u << H.fidx_bits) - 1u;
        for (uint64_t b = 0; b < H.fidx_nbuckets; ++b)
            for (int sl = 0; sl < kFidxEntries; ++sl) {
                const uint32_t v = e[b * kFidxBucketWords + sl];
                if (v && ((v & mask) == 0 || (v & mask) - 1u >= H.frec_units)) return "image: a flow index entry points outside the record array";
            }
        const uint32_t* u = (const uint32_t*)(img + H.frec_off);
        for (uint64_t i = 0; i < H.frec_units; ++i)
            if ((u[i * 8] & kSlotValid) && !res_ok(u[i * 8 + 2], u[i * 8 + 3])) {
                // the second unit of an IPv6 record holds address words, not a tag: only units an index entry names count
                bool named = false;
                for (uint64_t b = 0; b < (uint64_t)H.fidx_nbuckets * kFidxBucketWords && !named; ++b)
                    named = (b % kFidxBucketWords) < (uint64_t)kFidxEntries && e[b] && (e[b] & mask) - 1u == i;
                if (named) return "image: a flow record holds an unknown action or stage";
            }
    }

    // Everything the kernel follows as an index is checked here once, so that a damaged or hostile image
    // is refused on the host instead of faulting on the device.  Flow and L2 tables hold no indices.
    auto lpm_ok = [&](uint64_t root_off, uint64_t nodes_off, uint32_t nnodes, uint64_t tail_off, uint32_t ntails,
                      uint64_t nh_off) {
        auto value_ok = [&](uint32_t v) { return v == 0 || inside(nh_off + (uint64_t)(v - 1) * sizeof(LpmRes), sizeof(LpmRes)); };
        auto entry_ok = [&](uint32_t e) {
            if (e & kLpmChild) return (e & ~kLpmChild) < nnodes;
            if (e & kLpmTail) return (e & ~kLpmTail) < ntails;
            return value_ok(e);
        };
        const uint32_t* root = (const uint32_t*)(img + root_off);
        for (uint32_t i = 0; i < 65536; ++i)
            if (!entry_ok(root[i])) return false;
        const uint32_t* nodes = (const uint32_t*)(img + nodes_off);
        for (uint64_t i = 0; i < (uint64_t)nnodes * 256; ++i)
            if (!entry_ok(nodes[i])) return false;
        const LpmTail* tails = (const LpmTail*)(img + tail_off);
        for (uint32_t i = 0; i < ntails; ++i)
            if (tails[i].len < 1 || tails[i].len > 128 || tails[i].match == 0 || (tails[i].match & (kLpmChild | kLpmTail)) || (tails[i].fallback & (kLpmChild | kLpmTail)) ||
                !value_ok(tails[i].match) || !value_ok(tails[i].fallback))
                return false;
        return true;
    };
    if (!lpm_ok(H.lpm4_root_off, H.lpm4_nodes_off, H.lpm4_nnodes, H.lpm4_tail_off, H.lpm4_ntails, H.lpm4_nh_off) ||
        !lpm_ok(H.lpm6_root_off, H.lpm6_nodes_off, H.lpm6_nnodes, H.lpm6_tail_off, H.lpm6_ntails, H.lpm6_nh_off))
        return "image: an LPM entry points outside its table";
    if (H.acl_nrules) {
        const uint16_t* leaves = (const uint16_t*)(img + H.acl_leaf_off);
        auto cand_ok = [&](uint32_t c) { return c == kAclNone || (c & kAclRankMask) < H.acl_nrules; };
        auto field_ok = [&](uint32_t f) { return (f & 0x60u) == 0 && (f >> 7) < (uint32_t)kAclKeyRows; };
        auto entry_ok = [&](const AclEntry& e) {
            const uint32_t h[6] = {e.y & 0xFFFFu, e.y >> 16, e.z & 0xFFFFu, e.z >> 16, e.w & 0xFFFFu, e.w >> 16};
            if (e.x & kAclLeafBit) {
                for (uint32_t c : h)
                    if (!cand_ok(c)) return false;
                if (e.x & kAclMoreBit) {
                    const uint64_t off = e.x & kAclOffMask;
                    if (off >= H.acl_nleafwords || off + 1 + leaves[off] > H.acl_nleafwords) return false;
                    for (uint32_t i = 1; i <= leaves[off]; ++i)
                        if (!cand_ok(leaves[off + i]) || leaves[off + i] == kAclNone) return false;
                }
                return true;
            }
            uint64_t max_idx = 0;  // fields naming the zero row contribute a 0 bit wherever they stand
            for (uint32_t f : h) {
                if (!field_ok(f)) return false;
                max_idx = (max_idx << 1) | ((f >> 7) != 11u ? 1u : 0u);
            }
            return (uint64_t)e.x + max_idx < H.acl_nentries;
        };
        const AclEntry* nodes = (const AclEntry*)(img + H.acl_nodes_off);
        for (uint32_t t = 0; t < H.acl_ntrees; ++t)
            if (!entry_ok(H.acl_roots[t])) return "image: an ACL root entry is out of range";
        for (uint32_t i = 0; i < H.acl_nentries; ++i)
            if (!entry_ok(nodes[i])) return "image: an ACL tree entry is out of range";
    }
    return nullptr;
}

}  // namespace sfc
