// api.cu — the C ABI of libb200sfc.so (include/b200sfc.h): rule staging, image commit,
// kernel launches and the synthetic-traffic generator.  No CPU classification path exists in
// this library; without a CUDA device sfc_create fails.
#include <cuda.h>
#include <cuda_runtime.h>
#include <errno.h>

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <vector>

#include "../../include/b200sfc.h"
#include "classify.cuh"
#include "gen.h"
#include "tables.h"

using namespace sfc;

namespace {
constexpr int kTile = 256;
constexpr int kStages = 2;
#ifndef SFC_TILE_ACL
#define SFC_TILE_ACL 256
#endif
constexpr int kTileAcl = SFC_TILE_ACL;
#ifndef SFC_STAGES_ACL
#define SFC_STAGES_ACL 1
#endif
// The ACL variants are bound by table latency, not by the packet stream: one stage is enough (a slot is
// released as soon as its records sit in registers) and the shared memory it frees becomes L1 for the
// tree nodes and rule records (configs[3]: 63 -> 58 us per Mi packets).
constexpr int kStagesAcl = SFC_STAGES_ACL;
template <bool A>
constexpr int stages_for() { return A ? kStagesAcl : kStages; }
constexpr int kTicketRing = 256;     // a launch uses slot seq % 256 (so up to 256 launches may be in flight); slots are 32 bytes apart and self-resetting
constexpr int kTicketStride = 8;     // u32 words per slot: {draw counter, finished producers, pad}
constexpr size_t kHostChunk = 1u << 18;  // largest chunk of sfc_classify_host's pipeline (size of its device buffers)
constexpr size_t kHostChunkDefault = 1u << 18;
constexpr int kHostBufs = 4, kHostBufsDefault = 2;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
}  // namespace

struct sfc_handle {
    int device = 0;
    int sm_count = 0;
    std::string err;
    RuleStore rs;
    // committed image
    uint8_t* d_image = nullptr;
    size_t image_bytes = 0;
    uint64_t epoch = 0;
    DevTables T{};
    unsigned long long* d_counters = nullptr;
    std::atomic<uint64_t> launches{0};  // data-plane threads on different streams share the handle
    EncodeTiledFn encode = nullptr;
    int ctas_per_sm[40] = {0};  // 0-3 shipped variants, 4-15 tune variants (make tune), 16-19 multi-batch variants, 20-21 40-register variants
    bool low_regs = false;      // committed flow tables are too big for L2: use the 40-register variants
    int tune_tile = 0, tune_stages = 0;  // SFC_TUNE=tile,stages (experiments)
    int tune_ctas = 0, tune_carveout = -1;  // cap on resident CTAs per SM / shared-memory carveout percent
    uint32_t flow_first[2] = {0, 0};
    int probe_mode = -1;  // option "flow_probe_mode": -1 auto, else kProbe* bits
    int occ_mode = -1;    // option "low_regs": -1 by table size, 0 never, 1 always
    uint32_t* d_tickets = nullptr;  // ring of kTicketRing tile-ticket counters (scratch of the classify kernel)
    // sfc_classify_host staging
    cudaStream_t hs[kHostBufs] = {};
    cudaEvent_t ev[kHostBufs][2] = {};
    uint8_t* d_hdr[kHostBufs] = {};
    uint16_t* d_port[kHostBufs] = {};
    uint64_t* d_res[kHostBufs] = {};
    size_t host_chunk = kHostChunkDefault;  // option "host_chunk": packets per pipelined chunk
    int host_bufs = kHostBufsDefault;       // option "host_bufs": chunks in flight
    bool host_timing = false;               // option "host_timing": measure kernel time for sfc_last_kernel_ms
    uint64_t host_seq = 0;                  // chunks submitted so far; chunk k uses ring slot k % host_bufs
    bool host_busy[kHostBufs] = {};         // slot holds a chunk that has not been waited for
    uint64_t host_last[kHostBufs] = {};     // number of the chunk last submitted on the slot
    float host_ms_acc = 0.f;
    float last_kernel_ms = -1.f;
    // Host copy of the committed image (host_image below) + views of its flow tables: rule edits are applied
    // to the copy, every 32-byte sector they change is noted, and a commit ships exactly those sectors.
    FlowTable shadow4, shadow6;  // format 1 views into host_image
    IndexedFlows shadow2;        // format 2 view into host_image
    ImageHeader header{};        // header of the committed image (host copy)
    bool shadow_valid = false;
    uint8_t* h_patch = nullptr;  // pinned: [sector numbers u32 ...][sector data ...]
    uint8_t* d_patch = nullptr;
    size_t patch_cap = 0;
    cudaEvent_t patch_done = nullptr;
    cudaStream_t cs = nullptr;   // the library's own stream for sfc_commit
    uint64_t n_full_commits = 0, n_incremental_commits = 0, n_patched_buckets = 0;
    // `T` is what a launch copies; commits replace it.  Launches hold the lock shared from the copy until the
    // kernel is enqueued, a commit holds it exclusively while it swaps T — so once a commit has published, every
    // launch that can still see the old tables is already in a stream, and a device-wide wait retires them.
    std::shared_mutex tmu;
    // host copy of the committed image (empty after an import from device memory); for offline handles
    // (device_ordinal -1: rule staging and image building only, no CUDA calls) this IS the committed image
    std::vector<uint8_t> host_image;
    bool offline() const { return device < 0; }
};

namespace {

int fail(sfc_handle* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    return code;
}
#define CUDA_TRY(h, expr)                                                                        \
    do {                                                                                         \
        cudaError_t e_ = (expr);                                                                 \
        if (e_ != cudaSuccess) return fail(h, -EIO, "%s: %s", #expr, cudaGetErrorString(e_));    \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        if (dev < 0) return;
        // A thread that has made no runtime call yet reports device 0 as current but has no driver context
        // bound, and cuTensorMapEncodeTiled (a driver entry point) then fails with INVALID_CONTEXT: the
        // first call of every thread binds the primary context explicitly.
        static thread_local int bound = -1;
        cudaGetDevice(&prev);
        if (prev != dev || bound != dev) {
            cudaSetDevice(dev);
            bound = dev;
        }
        if (prev == dev) prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// Resolve the section offsets of an image header against the device address the image sits at.
void fill_tables(DevTables& T, const ImageHeader& H, const uint8_t* b, unsigned long long* counters, int probe_mode) {
    T.flow4 = (const uint32_t*)(b + H.flow4_off);
    T.flow6 = (const uint32_t*)(b + H.flow6_off);
    expand_seed(H.flow4_seed, T.S4, kFlow4SeedWords);
    expand_seed(H.flow6_seed, T.S6, kFlow6SeedWords);
    T.flow4_nbuckets = H.flow4_nbuckets;
    T.flow4_nhash = H.flow4_nhash;
    T.flow4_count = H.flow4_count;
    {   // probe strategy by where the table will live.  Measured on B200 (tools/tune.py): a table of
        // <= 128 KB stays in L1 and both probes at once win (14.8 vs 15.8 us per 1M packets); one that
        // stays in L2 next to the packet stream (measured up to 42 MB: 14.3 vs 15.5 us) still gains from
        // issuing both probes at once, without L1; beyond that every probe is a DRAM sector and the
        // second load only adds traffic (eager is 8-40% slower).
        uint64_t bytes = (uint64_t)H.flow4_nbuckets * H.flow4_slots * kFlow4BucketBytes;
        uint32_t mode = bytes <= (128u << 10) ? (kProbeEager | kProbeL1) : bytes <= kFlowL2ResidentBytes ? kProbeEager : 0u;
        if (probe_mode >= 0) mode = (uint32_t)probe_mode;
        T.flow4_mode = mode;
    }
    T.flow4_slots = H.flow4_slots;
    T.flow6_nbuckets = H.flow6_nbuckets;
    T.flow6_slots = H.flow6_slots;
    T.flow6_nhash = H.flow6_nhash;
    T.flow6_count = H.flow6_count;
    T.flow_format = H.flow_format;
    T.fidx = (const uint32_t*)(b + H.fidx_off);
    T.frec = (const uint32_t*)(b + H.frec_off);
    T.fidx_nbuckets = H.fidx_nbuckets;
    T.fidx_bits = H.fidx_bits;
    expand_seed(H.fidx_seed, T.S2, kFlow6SeedWords);
    if (H.flow_format == kFlowFormatIndexed && probe_mode < 0) {
        // index + records of a small rule set live in L1 (configs[1]: 6 KB + 32 KB)
        const uint64_t bytes = (uint64_t)H.fidx_nbuckets * 32 + (uint64_t)H.frec_used * 32;
        T.flow4_mode = bytes <= (96u << 10) ? kProbeL1 : 0u;
    }
    T.l2x = (const L2Entry*)(b + H.l2x_off);
    T.l2x_mask = H.l2x_mask;
    T.l2x_count = H.l2x_count;
    T.l2w = (const L2Wild*)(b + H.l2w_off);
    T.l2w_n = H.l2w_n;
    T.l2w_count = H.l2w_count;
    T.lpm4_root = (const uint32_t*)(b + H.lpm4_root_off);
    T.lpm4_nodes = (const uint32_t*)(b + H.lpm4_nodes_off);
    T.lpm4_nh = (const LpmRes*)(b + H.lpm4_nh_off);
    T.lpm4_tails = (const LpmTail*)(b + H.lpm4_tail_off);
    T.lpm4_count = H.lpm4_count;
    T.lpm6_root = (const uint32_t*)(b + H.lpm6_root_off);
    T.lpm6_nodes = (const uint32_t*)(b + H.lpm6_nodes_off);
    T.lpm6_nh = (const LpmRes*)(b + H.lpm6_nh_off);
    T.lpm6_tails = (const LpmTail*)(b + H.lpm6_tail_off);
    T.lpm6_count = H.lpm6_count;
    T.acl_nodes = (const uint4*)(b + H.acl_nodes_off);
    T.acl_leaves = (const uint16_t*)(b + H.acl_leaf_off);
    T.acl_rules = (const uint32_t*)(b + H.acl_rule_off);
    T.acl_res = (const uint32_t*)(b + H.acl_res_off);
    T.acl_nrules = H.acl_nrules;
    T.acl_ntrees = H.acl_ntrees;
    static_assert(sizeof(T.acl_roots) == sizeof(H.acl_roots), "root entries are copied verbatim");
    memcpy(T.acl_roots, H.acl_roots, sizeof(T.acl_roots));
    T.counters = counters;
}

// Which register budget the committed tables want (see launch_classify).
bool wants_low_regs(const sfc_handle* h, const ImageHeader& H) {
    // what the probes of one launch range over: flow slots, trie nodes and tail records
    uint64_t probed = (uint64_t)H.flow4_nbuckets * H.flow4_slots * kFlow4BucketBytes * (H.flow4_count ? 1 : 0) +
                      (uint64_t)H.flow6_nbuckets * H.flow6_slots * kFlow6BucketBytes * (H.flow6_count ? 1 : 0);
    if (H.flow_format == kFlowFormatIndexed) probed = (uint64_t)H.fidx_nbuckets * 32 + (uint64_t)H.frec_used * 32;
    if (H.lpm4_count) probed += (uint64_t)H.lpm4_nnodes * 1024 + (uint64_t)H.lpm4_ntails * sizeof(LpmTail);
    if (H.lpm6_count) probed += (uint64_t)H.lpm6_nnodes * 1024 + (uint64_t)H.lpm6_ntails * sizeof(LpmTail);
    // measured (profiles/r01_tuning.md): no difference around 20 MB, 5-10 % in favour of 40 registers from 30 MB up,
    // 7 % against when the table sits in L1
    return h->occ_mode < 0 ? probed > (24ull << 20) : h->occ_mode > 0;
}

// Make `d` (a device copy of an image whose header is H) the committed image: publish the new table
// descriptor, then wait until every launch that may still read the old image has finished, then free it.
int publish_image(sfc_handle* h, uint8_t* d, size_t len, const ImageHeader& H) {
    DevTables T{};
    fill_tables(T, H, d, h->d_counters, h->probe_mode);
    uint8_t* old = nullptr;
    {
        std::unique_lock<std::shared_mutex> lk(h->tmu);
        old = h->d_image;
        h->T = T;
        h->d_image = d;
        h->image_bytes = len;
        h->low_regs = wants_low_regs(h, H);
        h->epoch = H.epoch;
        h->flow_first[0] = H.flow_format == kFlowFormatIndexed ? H.fidx_first : H.flow4_first;
        h->flow_first[1] = H.flow6_first;
    }
    if (old) {
        // launches enqueued before the swap may still be reading `old`; none can be enqueued on it any more
        cudaError_t e = cudaDeviceSynchronize();
        cudaFree(old);
        if (e != cudaSuccess) return fail(h, -EIO, "cudaDeviceSynchronize: %s", cudaGetErrorString(e));
    }
    return 0;
}

int install_image_from_host(sfc_handle* h, const uint8_t* img, size_t len) {
    uint8_t* d = nullptr;
    CUDA_TRY(h, cudaMalloc(&d, len));
    cudaError_t e = cudaMemcpy(d, img, len, cudaMemcpyHostToDevice);  // returns when the copy has landed
    if (e != cudaSuccess) {
        cudaFree(d);
        return fail(h, -EIO, "image upload: %s", cudaGetErrorString(e));
    }
    ImageHeader H;
    memcpy(&H, img, sizeof(H));
    return publish_image(h, d, len, H);
}

template <int TILE, int STAGES, bool C, bool A, int OCC = 0, int FMT = 1>
int launch_variant(sfc_handle* h, int slot, const void* hdr, const uint16_t* in_port, uint64_t* result, size_t n,
                   uint32_t flags, cudaStream_t st) {
    auto kern = classify_kernel<TILE, STAGES, C, A, false, OCC, FMT>;
    constexpr int smem = SmemLayout<TILE, STAGES, A>::kBytes;
    if (!h->ctas_per_sm[slot]) {
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        if (h->tune_carveout >= 0)
            CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, h->tune_carveout));
        int per = 0;
        CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, kern, TILE + 32, smem));
        if (per < 1) return fail(h, -EIO, "classify kernel does not fit on an SM");
        if (h->tune_ctas > 0 && per > h->tune_ctas) per = h->tune_ctas;
        h->ctas_per_sm[slot] = per;
    }
    CUtensorMap map;
    cuuint64_t dims[2] = {16, (cuuint64_t)n};  // 16 x u32 per record, n records
    cuuint64_t strides[1] = {64};
    cuuint32_t box[2] = {16, (cuuint32_t)(TILE < 256 ? TILE : 256)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = h->encode(&map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(hdr), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(h, -EIO, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    uint64_t n_tiles = (n + TILE - 1) / TILE;
    // Grid policy (profiles/r01_tuning.md): a launch that runs alone wants every resident CTA it can
    // get; launches that overlap (SFC_F_OVERLAP) do better with ONE CTA per SM each, because the SM's
    // remaining room is what lets the next launches become resident early — several small grids in
    // flight beat one full grid that the next launch has to wait behind (12.3 vs 14.5 us per step).
    int per_sm = h->ctas_per_sm[slot];
    if ((flags & SFC_F_OVERLAP) && h->tune_ctas == 0) per_sm = 1;
    uint64_t grid = std::min<uint64_t>(n_tiles, (uint64_t)h->sm_count * per_sm);
    uint32_t* ticket = h->d_tickets + (h->launches.fetch_add(1) % kTicketRing) * kTicketStride;
    uint32_t overlap = (flags & SFC_F_OVERLAP) ? 1u : 0u;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(TILE + 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = overlap ? 1 : 0;
    CUDA_TRY(h, cudaLaunchKernelEx(&cfg, kern, map, in_port, result, (uint32_t)n, (uint32_t)n_tiles, ticket, overlap,
                                   h->T, NoBatches{}));
    return 0;
}

// One launch over the batches of `L` (already laid out: row0 / tile0 filled in, `rows` = extent of the
// tensor map in 64-byte records from L.base).
template <bool C, bool A, int OCC = 0, int FMT = 1>
int launch_multi(sfc_handle* h, int slot, const BatchList& L, uint64_t rows, uint32_t n_tiles, uint32_t flags,
                 cudaStream_t st) {
    auto kern = classify_kernel<kTile, stages_for<A>(), C, A, true, OCC, FMT>;
    constexpr int smem = SmemLayout<kTile, stages_for<A>(), A>::kBytes;
    if (!h->ctas_per_sm[slot]) {
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        int per = 0;
        CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, kern, kTile + 32, smem));
        if (per < 1) return fail(h, -EIO, "classify kernel does not fit on an SM");
        h->ctas_per_sm[slot] = per;
    }
    CUtensorMap map;
    cuuint64_t dims[2] = {16, (cuuint64_t)rows};
    cuuint64_t strides[1] = {64};
    cuuint32_t box[2] = {16, (cuuint32_t)(kTile < 256 ? kTile : 256)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = h->encode(&map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<uint8_t*>(L.base), dims, strides, box,
                           estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(h, -EIO, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    int per_sm = (flags & SFC_F_OVERLAP) ? 1 : h->ctas_per_sm[slot];
    uint64_t grid = std::min<uint64_t>(n_tiles, (uint64_t)h->sm_count * per_sm);
    uint32_t* ticket = h->d_tickets + (h->launches.fetch_add(1) % kTicketRing) * kTicketStride;
    uint32_t overlap = (flags & SFC_F_OVERLAP) ? 1u : 0u;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kTile + 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = overlap ? 1 : 0;
    CUDA_TRY(h, cudaLaunchKernelEx(&cfg, kern, map, (const uint16_t*)nullptr, (uint64_t*)nullptr, 0u, n_tiles, ticket,
                                   overlap, h->T, L));
    return 0;
}

int launch_classify(sfc_handle* h, const void* hdr, const uint16_t* in_port, uint64_t* result, size_t n,
                    uint32_t flags, cudaStream_t st) {
    if (((uintptr_t)hdr & 15) != 0) return fail(h, -EINVAL, "sfc_classify: hdr must be 16-byte aligned");
    if (n > (1ull << 31)) return fail(h, -EINVAL, "sfc_classify: at most 2^31 packets per call");
    // held from the choice of the kernel variant until the launch is in its stream (see sfc_handle::tmu)
    std::shared_lock<std::shared_mutex> lk(h->tmu);
    if (!h->d_image) return fail(h, -EINVAL, "sfc_classify: no committed table image (call sfc_commit)");
    bool c = flags & SFC_F_COUNTERS, a = h->T.acl_nrules > 0;
#ifdef SFC_TUNE_BUILD  // experiment hook (make tune): SFC_TUNE=tile,stages or options tune_tile / tune_stages
    if (!c && !a && h->tune_tile) {
#define SFC_TUNE_CASE(slot, T_, S_) \
    if (h->tune_tile == T_ && h->tune_stages == S_) return launch_variant<T_, S_, false, false>(h, slot, hdr, in_port, result, n, flags, st);
        SFC_TUNE_CASE(4, 64, 4)
        SFC_TUNE_CASE(5, 128, 2)
        SFC_TUNE_CASE(6, 128, 3)
        SFC_TUNE_CASE(7, 256, 2)
        SFC_TUNE_CASE(8, 256, 3)
        SFC_TUNE_CASE(9, 256, 4)
        SFC_TUNE_CASE(10, 512, 2)
        SFC_TUNE_CASE(11, 512, 3)
        SFC_TUNE_CASE(12, 768, 2)
        SFC_TUNE_CASE(13, 256, 6)
        SFC_TUNE_CASE(14, 256, 8)
        SFC_TUNE_CASE(15, 128, 4)
#undef SFC_TUNE_CASE
    }
#endif
    // one compiled kernel per (counters, ACL, register budget, flow table format); slot = index into ctas_per_sm
    const bool low = h->low_regs && !a;  // probes wait on L2 / DRAM: five CTAs per SM at 40 registers (cfg3: 20.9 vs 22.3 us per Mi)
#define SFC_GO(slot, C_, A_, OCC_, FMT_)                                                                              \
    return launch_variant<(A_) ? kTileAcl : kTile, (A_) ? kStagesAcl : kStages, C_, A_, OCC_, FMT_>(h, slot, hdr, in_port, \
                                                                                                  result, n, flags, st)
    if (h->T.flow_format == kFlowFormatIndexed) {
        if (c && a) SFC_GO(27, true, true, 0, 2);
        if (low && c) SFC_GO(29, true, false, 5, 2);
        if (low) SFC_GO(28, false, false, 5, 2);
        if (c) SFC_GO(26, true, false, 0, 2);
        if (a) SFC_GO(25, false, true, 0, 2);
        SFC_GO(24, false, false, 0, 2);
    }
    if (c && a) SFC_GO(3, true, true, 0, 1);
    if (low && c) SFC_GO(21, true, false, 5, 1);
    if (low) SFC_GO(20, false, false, 5, 1);
    if (c) SFC_GO(2, true, false, 0, 1);
    if (a) SFC_GO(1, false, true, 0, 1);
    SFC_GO(0, false, false, 0, 1);
#undef SFC_GO
}

// Incremental commit: write `n` 32-byte sectors into the image in place, one thread per sector and one
// 32-byte store per thread, so that a lookup running beside it sees each sector either old or new.
__global__ void patch_sectors_kernel(uint8_t* image, const uint32_t* __restrict__ sector, const uint4* __restrict__ data,
                                     uint32_t n) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const uint4 a = data[2 * t], b = data[2 * t + 1];
    asm volatile("st.global.v8.u32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(image + (size_t)sector[t] * 32), "r"(a.x),
                 "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w)
                 : "memory");
}

__global__ void gen_kernel(sfc_gen_spec spec, uint64_t first, uint64_t n, uint8_t* hdr, uint16_t* in_port) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint8_t h[64];
    uint16_t port;
    gen_packet(spec, first + i, h, port);
    uint4* dst = reinterpret_cast<uint4*>(hdr + i * 64);
    const uint4* src = reinterpret_cast<const uint4*>(h);
#pragma unroll
    for (int k = 0; k < 4; ++k) dst[k] = src[k];
    in_port[i] = port;
}

}  // namespace

// =================================================================== life cycle ==========
extern "C" {

const char* sfc_version(void) { return "b200sfc 0.9 (image v12)"; }

int sfc_create(int device_ordinal, sfc_handle** out) {
    if (!out) return -EINVAL;
    *out = nullptr;
    if (device_ordinal == -1) {  // offline: stage rules and build images, nothing else
        sfc_handle* h = new sfc_handle();
        h->device = -1;
        *out = h;
        return 0;
    }
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return -ENODEV;
    if (device_ordinal < 0 || device_ordinal >= count) return -ENODEV;
    sfc_handle* h = new sfc_handle();
    h->device = device_ordinal;
    DeviceGuard g(device_ordinal);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device_ordinal) != cudaSuccess || prop.major < 10) {
        delete h;
        return -ENODEV;  // sm_100a code only
    }
    h->sm_count = prop.multiProcessorCount;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) {
        delete h;
        return -ENOSYS;
    }
    h->encode = (EncodeTiledFn)fn;
#ifdef SFC_TUNE_BUILD
    if (const char* tune = getenv("SFC_TUNE")) sscanf(tune, "%d,%d", &h->tune_tile, &h->tune_stages);
#endif
    if (cudaMalloc(&h->d_counters, SFC_CNT_TOTAL * sizeof(uint64_t)) != cudaSuccess ||
        cudaMemset(h->d_counters, 0, SFC_CNT_TOTAL * sizeof(uint64_t)) != cudaSuccess ||
        cudaMalloc(&h->d_tickets, kTicketRing * kTicketStride * sizeof(uint32_t)) != cudaSuccess ||
        cudaMemset(h->d_tickets, 0, kTicketRing * kTicketStride * sizeof(uint32_t)) != cudaSuccess) {
        delete h;
        return -ENOMEM;
    }
    *out = h;
    return 0;
}

int sfc_destroy(sfc_handle* h) {
    if (!h) return -EINVAL;
    if (h->offline()) {
        delete h;
        return 0;
    }
    DeviceGuard g(h->device);
    cudaDeviceSynchronize();
    if (h->d_image) cudaFree(h->d_image);
    if (h->d_counters) cudaFree(h->d_counters);
    if (h->d_tickets) cudaFree(h->d_tickets);
    if (h->h_patch) cudaFreeHost(h->h_patch);
    if (h->d_patch) cudaFree(h->d_patch);
    if (h->patch_done) cudaEventDestroy(h->patch_done);
    if (h->cs) cudaStreamDestroy(h->cs);
    for (int i = 0; i < kHostBufs; ++i) {
        if (h->d_hdr[i]) cudaFree(h->d_hdr[i]);
        if (h->d_port[i]) cudaFree(h->d_port[i]);
        if (h->d_res[i]) cudaFree(h->d_res[i]);
        if (h->hs[i]) cudaStreamDestroy(h->hs[i]);
        for (int j = 0; j < 2; ++j)
            if (h->ev[i][j]) cudaEventDestroy(h->ev[i][j]);
    }
    delete h;
    return 0;
}

const char* sfc_last_error(const sfc_handle* h) { return h ? h->err.c_str() : "null handle"; }

int sfc_set_option(sfc_handle* h, const char* name, long value) {
    if (!h || !name) return -EINVAL;
    if (!strcmp(name, "flow_load_pct")) {
        if (value != 0 && (value < 5 || value > 90)) return fail(h, -EINVAL, "flow_load_pct must be 0 (the format's default) or 5..90");
        h->rs.other_dirty = true;
        h->rs.flow_load_pct = (int)value;
        return 0;
    }
    if (!strcmp(name, "flow_hashes")) {
        if (value < 2 || value > kMaxHashes) return fail(h, -EINVAL, "flow_hashes must be 2 or 3");
        h->rs.other_dirty = true;
        h->rs.flow_hashes = (int)value;
        return 0;
    }
    if (!strcmp(name, "host_chunk")) {  // packets per chunk of sfc_classify_host's copy/classify/copy pipeline
        if (value < 1024 || (size_t)value > kHostChunk) return fail(h, -EINVAL, "host_chunk must be 1024..%zu", kHostChunk);
        h->host_chunk = (size_t)value;
        return 0;
    }
    if (!strcmp(name, "host_timing")) {
        h->host_timing = value != 0;
        return 0;
    }
    if (!strcmp(name, "host_bufs")) {
        if (value < 2 || value > kHostBufs) return fail(h, -EINVAL, "host_bufs must be 2..%d", kHostBufs);
        if (!h->offline()) sfc_classify_host_wait(h, ~0ull);  // the slot of a chunk number changes with the ring size
        h->host_bufs = (int)value;
        return 0;
    }
    if (!strcmp(name, "low_regs")) {  // which register budget the non-ACL kernels run with: -1 by table size, 0 / 1 forced
        if (value < -1 || value > 1) return fail(h, -EINVAL, "low_regs must be -1, 0 or 1");
        h->occ_mode = (int)value;
        h->low_regs = value < 0 ? h->low_regs : value > 0;
        h->rs.other_dirty = true;  // the size rule is re-evaluated at the next full commit
        return 0;
    }
    if (!strcmp(name, "flow_slots")) {  // slots per flow bucket: 0 / 1 = one (default), 2 = dense two-slot buckets
        if (value < 0 || value > 2) return fail(h, -EINVAL, "flow_slots must be 0, 1 or 2");
        h->rs.other_dirty = true;
        h->rs.flow_slots = (int)value;
        return 0;
    }
    if (!strcmp(name, "flow_probe_mode")) {  // -1 auto, bit0 eager dual probe, bit1 probes may live in L1
        if (value < -1 || value > 3) return fail(h, -EINVAL, "flow_probe_mode must be -1..3");
        h->probe_mode = (int)value;
        std::unique_lock<std::shared_mutex> lk(h->tmu);
        if (h->d_image) h->T.flow4_mode = value >= 0 ? (uint32_t)value : h->T.flow4_mode;
        return 0;
    }
    if (!strcmp(name, "flow_format")) {  // 0 by size, 1 one-sector cuckoo tables per family, 2 fingerprint index + records
        if (value < 0 || value > 2) return fail(h, -EINVAL, "flow_format must be 0, 1 or 2");
        h->rs.other_dirty = true;
        h->rs.flow_format = (int)value;
        return 0;
    }
#ifdef SFC_TUNE_BUILD
    if (!strcmp(name, "tune_tile")) {  // experiment hooks: pick a compiled (tile, stages) variant
        h->tune_tile = (int)value;
        return 0;
    }
    if (!strcmp(name, "tune_stages")) {
        h->tune_stages = (int)value;
        return 0;
    }
    if (!strcmp(name, "tune_ctas") || !strcmp(name, "tune_carveout")) {
        (name[5] == 'c' && name[6] == 't' ? h->tune_ctas : h->tune_carveout) = (int)value;
        for (int& c : h->ctas_per_sm) c = 0;  // re-derive occupancy on the next launch
        return 0;
    }
#endif
    return fail(h, -ENOENT, "unknown option '%s'", name);
}

// =================================================================== ports ===============
int sfc_port_add(sfc_handle* h, const char* name, uint16_t* port_id) {
    if (!h || !name || !*name) return h ? fail(h, -EINVAL, "sfc_port_add: empty name") : -EINVAL;
    auto it = h->rs.port_by_name.find(name);
    if (it != h->rs.port_by_name.end()) {
        if (port_id) *port_id = it->second;
        return 0;
    }
    size_t id = 0;
    while (id < h->rs.port_names.size() && !h->rs.port_names[id].empty()) ++id;
    if (id >= SFC_MAX_PORTS) return fail(h, -ENOSPC, "sfc_port_add: all %u port ids are in use", SFC_MAX_PORTS);
    if (id == h->rs.port_names.size()) h->rs.port_names.emplace_back();
    h->rs.port_names[id] = name;
    h->rs.port_by_name[name] = (uint16_t)id;
    if (port_id) *port_id = (uint16_t)id;
    return 0;
}

int sfc_port_del(sfc_handle* h, const char* name) {
    if (!h || !name) return -EINVAL;
    auto it = h->rs.port_by_name.find(name);
    if (it == h->rs.port_by_name.end()) return fail(h, -ENOENT, "sfc_port_del: no port named '%s'", name);
    // The id goes back into the pool (sfc_port_add hands out the lowest free one), so steering rules that name
    // it must not outlive it: with OVS a flow that names a removed port stops matching; here it is dropped.
    const uint16_t id = it->second;
    auto& v = h->rs.l2;
    const size_t before = v.size();
    v.erase(std::remove_if(v.begin(), v.end(), [&](const L2Rule& e) { return e.in_port == id || e.out_port == id; }), v.end());
    if (v.size() != before) h->rs.l2_dirty = true;
    h->rs.port_names[id].clear();
    h->rs.port_by_name.erase(it);
    return 0;
}

int sfc_port_lookup(const sfc_handle* h, const char* name, uint16_t* port_id) {
    if (!h || !name) return -EINVAL;
    auto it = h->rs.port_by_name.find(name);
    if (it == h->rs.port_by_name.end()) return -ENOENT;
    if (port_id) *port_id = it->second;
    return 0;
}

long sfc_port_list(const sfc_handle* h, char* buf, size_t len) {
    if (!h) return -EINVAL;
    std::string s;
    for (auto& n : h->rs.port_names)
        if (!n.empty()) {
            s += n;
            s += '\n';
        }
    if (buf && len) {
        size_t c = std::min(len - 1, s.size());
        memcpy(buf, s.data(), c);
        buf[c] = 0;
    }
    return (long)s.size() + 1;
}

// =================================================================== L2 steering =========
int sfc_l2_rule_add(sfc_handle* h, uint16_t in_port, uint16_t out_port, const uint8_t* dst_mac, uint32_t* rule_id) {
    if (!h) return -EINVAL;
    h->rs.l2_dirty = true;  // the next commit rebuilds the idle copy of the L2 tables, not the image
    if (in_port >= SFC_MAX_PORTS || out_port >= SFC_MAX_PORTS)
        return fail(h, -EINVAL, "sfc_l2_rule_add: port id out of range");
    L2Rule r{};
    r.in_port = in_port;
    r.out_port = out_port;
    r.has_mac = dst_mac != nullptr;
    if (dst_mac) memcpy(r.mac, dst_mac, 6);
    // priorities exactly as ovs-dp/ovsdp.go:120,127,135 writes them
    r.priority = !dst_mac ? 10u : (in_port == out_port ? 100u : 32768u);
    for (auto& e : h->rs.l2)
        if (e.in_port == r.in_port && e.has_mac == r.has_mac && e.priority == r.priority &&
            (!r.has_mac || !memcmp(e.mac, r.mac, 6))) {
            e.out_port = out_port;  // add-flow on an identical match replaces the actions
            if (rule_id) *rule_id = e.id;
            return 0;
        }
    if (h->rs.next_l2_id >= (1u << 28)) return fail(h, -ENOSPC, "sfc_l2_rule_add: rule ids exhausted");
    r.id = h->rs.next_l2_id++;
    h->rs.l2.push_back(r);
    if (rule_id) *rule_id = r.id;
    return 0;
}

int sfc_l2_rule_del(sfc_handle* h, uint16_t in_port, const uint8_t* dst_mac) {
    if (!h) return -EINVAL;
    h->rs.l2_dirty = true;  // the next commit rebuilds the idle copy of the L2 tables, not the image
    auto& v = h->rs.l2;
    v.erase(std::remove_if(v.begin(), v.end(),
                           [&](const L2Rule& e) {
                               if (e.in_port != in_port) return false;
                               if (!dst_mac) return true;  // loose delete: every flow with this in_port
                               return e.has_mac && !memcmp(e.mac, dst_mac, 6);
                           }),
            v.end());
    return 0;
}

int sfc_chain_set(sfc_handle* h, const uint16_t* nf_in_ports, const uint16_t* nf_out_ports, size_t n_nfs) {
    if (!h) return -EINVAL;
    h->rs.other_dirty = true;  // not a flow edit: the next commit rebuilds the image
    if (n_nfs > SFC_MAX_NFS) return fail(h, -EINVAL, "sfc_chain_set: at most %u network functions", SFC_MAX_NFS);
    if (n_nfs && (!nf_in_ports || !nf_out_ports)) return fail(h, -EINVAL, "sfc_chain_set: null port array");
    for (size_t i = 0; i < n_nfs; ++i)
        if (nf_in_ports[i] >= SFC_MAX_PORTS || nf_out_ports[i] >= SFC_MAX_PORTS)
            return fail(h, -EINVAL, "sfc_chain_set: port id out of range");
    h->rs.nf_in.assign(nf_in_ports, nf_in_ports + n_nfs);
    h->rs.nf_out.assign(nf_out_ports, nf_out_ports + n_nfs);
    return 0;
}

// =================================================================== flows ===============
int sfc_flow_add_bulk(sfc_handle* h, const sfc_flow_key* keys, const sfc_flow_action* acts, size_t n,
                      uint32_t* first_id) {
    if (!h || (n && (!keys || !acts))) return -EINVAL;
    RuleStore& rs = h->rs;
    if ((uint64_t)rs.next_flow_id + n > SFC_MAX_FLOWS) return fail(h, -ENOSPC, "flow ids exhausted");
    size_t before = rs.flows.size(), log_before = rs.flow_log.size();
    const bool log_it = !rs.other_dirty && rs.flow_log.size() + n <= (1u << 20);
    uint32_t id0 = rs.next_flow_id;
    if (n > 4096) rs.flows.reserve(before + n);
    std::string err;
    for (size_t i = 0; i < n; ++i) {
        FlowRec f;
        int code = 0;
        if (!flow_from_key(keys[i], acts[i], f, err)) code = -EINVAL;
        else if (rs.flow_index.find(rs.flows, f) >= 0) {
            err = "flow already present";
            code = -EEXIST;
        }
        if (code) {  // roll the whole batch back
            rs.flows.resize(before);
            rs.flow_index.rebuild(rs.flows, before);
            rs.flow_log.resize(log_before);
            return fail(h, code, "sfc_flow_add: entry %zu: %s", i, err.c_str());
        }
        f.id = id0 + (uint32_t)i;
        rs.flows.push_back(f);
        rs.flow_index.insert(rs.flows, (uint32_t)(rs.flows.size() - 1));
        if (log_it) rs.flow_log.push_back(FlowChange{true, f});
    }
    if (!log_it) rs.other_dirty = true;  // a bulk load this large is cheaper as a rebuild
    rs.next_flow_id = id0 + (uint32_t)n;
    if (first_id) *first_id = id0;
    return 0;
}

int sfc_flow_add(sfc_handle* h, const sfc_flow_key* key, const sfc_flow_action* act, uint32_t* flow_id) {
    return sfc_flow_add_bulk(h, key, act, 1, flow_id);
}

int sfc_flow_del(sfc_handle* h, const sfc_flow_key* key) {
    if (!h || !key) return -EINVAL;
    FlowRec f;
    std::string err;
    sfc_flow_action a{SFC_ACT_TO_CP, 0, 0};
    if (!flow_from_key(*key, a, f, err)) return fail(h, -EINVAL, "sfc_flow_del: %s", err.c_str());
    int64_t pos = h->rs.flow_index.find(h->rs.flows, f);
    if (pos < 0) return fail(h, -ENOENT, "sfc_flow_del: no such flow");
    if (!h->rs.other_dirty) h->rs.flow_log.push_back(FlowChange{false, h->rs.flows[pos]});
    h->rs.flow_index.erase(h->rs.flows, (uint32_t)pos);
    return 0;
}

// =================================================================== LPM =================
static int lpm_words(sfc_handle* h, int af, const uint8_t* prefix, int len, uint32_t* w) {
    if (af != 4 && af != 6) return fail(h, -EINVAL, "lpm: af must be 4 or 6");
    if (!prefix || len < 0 || len > (af == 4 ? 32 : 128)) return fail(h, -EINVAL, "lpm: bad prefix length");
    for (int k = 0; k < 4; ++k) {
        uint32_t v = 0;
        if (af == 6 || k == 0)
            v = ((uint32_t)prefix[4 * k] << 24) | ((uint32_t)prefix[4 * k + 1] << 16) |
                ((uint32_t)prefix[4 * k + 2] << 8) | prefix[4 * k + 3];
        w[k] = v & mask32(len - 32 * k);
    }
    return 0;
}

int sfc_lpm_add(sfc_handle* h, int af, const uint8_t* prefix, int len, uint16_t out_port, uint32_t* prefix_id) {
    if (!h) return -EINVAL;
    h->rs.other_dirty = true;  // not a flow edit: the next commit rebuilds the image
    LpmRec r{};
    int rc = lpm_words(h, af, prefix, len, r.w);
    if (rc) return rc;
    if (out_port >= SFC_MAX_PORTS) return fail(h, -EINVAL, "sfc_lpm_add: out_port out of range");
    auto& v = af == 4 ? h->rs.lpm4 : h->rs.lpm6;
    RuleStore::LpmKey key = {(uint32_t)len, r.w[0], r.w[1], r.w[2], r.w[3]};
    auto it = v.find(key);
    if (it != v.end()) {  // same prefix again: replace the next hop, keep the id
        it->second.out_port = out_port;
        if (prefix_id) *prefix_id = it->second.id;
        return 0;
    }
    if (h->rs.next_lpm_id >= (1u << 28)) return fail(h, -ENOSPC, "sfc_lpm_add: prefix ids exhausted");
    r.len = len;
    r.out_port = out_port;
    r.id = h->rs.next_lpm_id++;
    v[key] = r;
    if (prefix_id) *prefix_id = r.id;
    return 0;
}

int sfc_lpm_del(sfc_handle* h, int af, const uint8_t* prefix, int len) {
    if (!h) return -EINVAL;
    h->rs.other_dirty = true;  // not a flow edit: the next commit rebuilds the image
    uint32_t w[4];
    int rc = lpm_words(h, af, prefix, len, w);
    if (rc) return rc;
    auto& v = af == 4 ? h->rs.lpm4 : h->rs.lpm6;
    RuleStore::LpmKey key = {(uint32_t)len, w[0], w[1], w[2], w[3]};
    if (!v.erase(key)) return fail(h, -ENOENT, "sfc_lpm_del: no such prefix");
    return 0;
}

// =================================================================== ACL =================
int sfc_acl_add(sfc_handle* h, const sfc_acl_rule* rule, uint32_t* rule_id) {
    if (!h || !rule) return -EINVAL;
    h->rs.other_dirty = true;
    if (h->rs.acl.size() >= SFC_MAX_ACL_RULES) return fail(h, -ENOSPC, "sfc_acl_add: table full");
    if (h->rs.next_acl_id >= SFC_MAX_ACL_RULES) return fail(h, -ENOSPC, "sfc_acl_add: rule ids exhausted (sfc_clear)");
    if (rule->ip_ver != 0 && rule->ip_ver != 4 && rule->ip_ver != 6) return fail(h, -EINVAL, "sfc_acl_add: ip_ver");
    if (rule->verdict > SFC_ACL_DENY) return fail(h, -EINVAL, "sfc_acl_add: verdict");
    for (int i = 0; i < 16; ++i) {
        bool m = rule->src_mask[i] || rule->dst_mask[i];
        if (m && rule->ip_ver == 0) return fail(h, -EINVAL, "sfc_acl_add: address mask needs ip_ver 4 or 6");
        if (m && rule->ip_ver == 4 && i >= 4) return fail(h, -EINVAL, "sfc_acl_add: IPv4 mask bytes 4..15 must be 0");
    }
    AclRec a;
    a.r = *rule;
    a.id = h->rs.next_acl_id++;
    h->rs.acl.push_back(a);
    if (rule_id) *rule_id = a.id;
    return 0;
}

int sfc_acl_del(sfc_handle* h, uint32_t rule_id) {
    if (!h) return -EINVAL;
    h->rs.other_dirty = true;  // not a flow edit: the next commit rebuilds the image
    for (size_t i = 0; i < h->rs.acl.size(); ++i)
        if (h->rs.acl[i].id == rule_id) {
            h->rs.acl.erase(h->rs.acl.begin() + i);
            return 0;
        }
    return fail(h, -ENOENT, "sfc_acl_del: no rule with id %u", rule_id);
}

int sfc_clear(sfc_handle* h) {
    if (!h) return -EINVAL;
    h->rs.clear_rules();
    return 0;
}

// =================================================================== commit / image ======
}  // extern "C" (reopened below)

namespace {

// Point the shadow views at the host copy of the committed image so that later edits can be applied
// sector by sector.
void adopt_shadows(sfc_handle* h) {
    uint8_t* base = h->host_image.data();
    memcpy(&h->header, base, sizeof(ImageHeader));
    const ImageHeader& H = h->header;
    auto take = [&](FlowTable& t, bool v6, uint64_t off, uint32_t nb, uint32_t slots, uint32_t nhash, uint64_t seed) {
        t = FlowTable();
        t.v6 = v6;
        t.tab = (uint32_t*)(base + off);
        t.nbuckets = nb;
        t.slots = (int)slots;
        t.nhash = (int)nhash;
        expand_seed(seed, t.S, v6 ? kFlow6SeedWords : kFlow4SeedWords);
    };
    take(h->shadow4, false, H.flow4_off, H.flow4_nbuckets, H.flow4_slots, H.flow4_nhash, H.flow4_seed);
    take(h->shadow6, true, H.flow6_off, H.flow6_nbuckets, H.flow6_slots, H.flow6_nhash, H.flow6_seed);
    IndexedFlows& x = h->shadow2;
    x = IndexedFlows();
    if (H.flow_format == kFlowFormatIndexed) {
        x.idx = (uint32_t*)(base + H.fidx_off);
        x.rec = (uint32_t*)(base + H.frec_off);
        x.nbuckets = H.fidx_nbuckets;
        x.bits = H.fidx_bits;
        x.units = H.frec_units;
        x.used = H.frec_used;
        x.idx_sector0 = (uint32_t)(H.fidx_off / 32);
        x.rec_sector0 = (uint32_t)(H.frec_off / 32);
        x.count = H.flow4_count + H.flow6_count;
        x.second = x.count - H.fidx_first;
        expand_seed(H.fidx_seed, x.S, kFlow6SeedWords);
    }
    h->shadow_valid = true;
}

int full_commit(sfc_handle* h) {
    std::vector<uint8_t> img;
    std::string err;
    if (!build_image(h->rs, h->epoch + 1, img, err)) return fail(h, -ENOSPC, "sfc_commit: %s", err.c_str());
    if (!h->offline()) {
        // the device first: if this fails the old image, its host copy and the edit log all stay as they were
        int rc = install_image_from_host(h, img.data(), img.size());
        if (rc) return rc;
    } else {
        h->image_bytes = img.size();
        ++h->epoch;
    }
    h->host_image.swap(img);
    adopt_shadows(h);
    h->rs.flow_log.clear();
    h->rs.l2_dirty = false;
    h->rs.other_dirty = false;
    ++h->n_full_commits;
    return 0;
}

// Apply the logged flow edits and a changed L2 rule set to the host copy of the image and ship only the
// sectors they changed.  Returns 1 if done, 0 if a full rebuild is needed, < 0 on error.
// Whenever 0 or an error is returned after the host copy was touched, shadow_valid is false: the next commit
// rebuilds from the rule store (the source of truth), so no edit is ever applied twice.
int incremental_commit(sfc_handle* h, cudaStream_t st, bool async) {
    RuleStore& rs = h->rs;
    if (rs.other_dirty || !h->shadow_valid || h->host_image.empty() || (!h->offline() && !h->d_image)) return 0;
    ImageHeader& H = h->header;
    const bool indexed = H.flow_format == kFlowFormatIndexed;
    long c4 = H.flow4_count, c6 = H.flow6_count, adds = 0;
    for (auto& ch : rs.flow_log) {
        (ch.rec.ip_ver == 6 ? c6 : c4) += ch.add ? 1 : -1;
        adds += ch.add;
    }
    if (indexed) {
        if (adds > h->shadow2.room()) return 0;
    } else {
        auto room = [&](const FlowTable& t) {  // flows the table may hold before it has to be rebuilt larger
            const long max_load = t.slots == 2 ? 85 : (rs.flow_hashes == 2 ? 45 : 85);
            return (long)t.nbuckets * t.slots * max_load / 100;
        };
        if (c4 > room(h->shadow4) || c6 > room(h->shadow6)) return 0;
    }
    std::vector<uint32_t> dirty;  // image sector numbers
    auto bail = [&](int rc) {
        h->shadow_valid = false;
        return rc;
    };
    // ---- flows ----
    uint32_t rec[kFlow6BucketWords];
    std::vector<uint32_t> slots;
    for (auto& ch : rs.flow_log) {
        const FlowRec& f = ch.rec;
        const uint32_t nf = f.action == SFC_ACT_FORWARD ? nf_of_port(rs, f.out_port) : (uint32_t)SFC_NF_NONE;
        bool ok;
        if (indexed) {
            ok = ch.add ? h->shadow2.insert(f, nf, &dirty) : h->shadow2.erase(f, &dirty);
        } else {
            const bool six = f.ip_ver == 6;
            FlowTable& t = six ? h->shadow6 : h->shadow4;
            FlowTable::encode(f, nf, rec);
            slots.clear();
            ok = ch.add ? t.insert(rec, &slots) : t.erase(rec, &slots);
            const uint64_t off = six ? H.flow6_off : H.flow4_off;
            for (uint32_t sl : slots) {  // an IPv4 slot is one sector, an IPv6 slot two
                const uint32_t s0 = (uint32_t)(off / 32) + sl * (six ? 2u : 1u);
                dirty.push_back(s0);
                if (six) dirty.push_back(s0 + 1);
            }
        }
        // A failed insertion walk leaves one flow homeless: fall back to a rebuild from the rule store.
        // Nothing has reached the device yet.
        if (!ok) return bail(0);
    }
    H.flow4_count = (uint32_t)c4;
    H.flow6_count = (uint32_t)c6;
    if (indexed) {
        H.frec_used = h->shadow2.used;
        H.fidx_first = h->shadow2.count - h->shadow2.second;
    }
    H.epoch = h->epoch + 1;
    memcpy(h->host_image.data(), &H, sizeof(ImageHeader));
    // ---- L2 steering: rebuilt into the idle copy, which becomes the active one ----
    if (rs.l2_dirty) {
        if (!h->offline() && !async) {
            // the idle copy was the active one before the previous commit: launches from back then must be gone
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) return bail(fail(h, -EIO, "cudaDeviceSynchronize: %s", cudaGetErrorString(e)));
        }
        if (!rebuild_l2_copy(rs, h->host_image.data(), 1 - (int)H.l2_active, &dirty)) return bail(0);
        memcpy(&H, h->host_image.data(), sizeof(ImageHeader));
    }
    for (uint32_t sct = 0; sct < kImageHeaderBytes / 32; ++sct) dirty.push_back(sct);
    std::sort(dirty.begin(), dirty.end());
    dirty.erase(std::unique(dirty.begin(), dirty.end()), dirty.end());
    if (!h->offline()) {
        // staging layout: [sector numbers][pad to 32][sector data]
        const size_t n = dirty.size(), o_data = (n * 4 + 31) & ~(size_t)31, need = o_data + n * 32;
        auto cuda_bail = [&](cudaError_t e, const char* what) { return bail(fail(h, -EIO, "%s: %s", what, cudaGetErrorString(e))); };
        cudaError_t e;
        if (!h->patch_done) {
            if ((e = cudaEventCreateWithFlags(&h->patch_done, cudaEventDisableTiming)) != cudaSuccess) return cuda_bail(e, "cudaEventCreate");
        } else if ((e = cudaEventSynchronize(h->patch_done)) != cudaSuccess) {  // the previous patch has left the staging buffer
            return cuda_bail(e, "cudaEventSynchronize");
        }
        if (!async && !h->cs && (e = cudaStreamCreateWithFlags(&h->cs, cudaStreamNonBlocking)) != cudaSuccess) return cuda_bail(e, "cudaStreamCreate");
        if (!async) st = h->cs;
        if (need > h->patch_cap) {
            if (h->h_patch) cudaFreeHost(h->h_patch);
            if (h->d_patch) cudaFree(h->d_patch);
            h->h_patch = h->d_patch = nullptr;
            h->patch_cap = 0;
            const size_t cap = std::max<size_t>(need * 2, 1 << 20);
            if ((e = cudaMallocHost(&h->h_patch, cap)) != cudaSuccess) return cuda_bail(e, "cudaMallocHost");
            if ((e = cudaMalloc(&h->d_patch, cap)) != cudaSuccess) return cuda_bail(e, "cudaMalloc");
            h->patch_cap = cap;
        }
        uint8_t* p = h->h_patch;
        memcpy(p, dirty.data(), n * 4);
        for (size_t i = 0; i < n; ++i) memcpy(p + o_data + i * 32, h->host_image.data() + (size_t)dirty[i] * 32, 32);
        if ((e = cudaMemcpyAsync(h->d_patch, p, need, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_bail(e, "cudaMemcpyAsync");
        patch_sectors_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(h->d_image, (const uint32_t*)h->d_patch,
                                                                           (const uint4*)(h->d_patch + o_data), (uint32_t)n);
        ++h->launches;
        if ((e = cudaGetLastError()) != cudaSuccess) return cuda_bail(e, "patch kernel");
        if ((e = cudaEventRecord(h->patch_done, st)) != cudaSuccess) return cuda_bail(e, "cudaEventRecord");
        // sfc_commit: the sectors have landed before the new descriptor is published, so launches on ANY stream
        // see a finished table.  sfc_commit_stream: published at once; launches on `st` are ordered behind the patch.
        if (!async && (e = cudaStreamSynchronize(st)) != cudaSuccess) return cuda_bail(e, "cudaStreamSynchronize");
        DevTables T{};
        fill_tables(T, H, h->d_image, h->d_counters, h->probe_mode);
        std::unique_lock<std::shared_mutex> lk(h->tmu);
        h->T = T;
        h->flow_first[0] = indexed ? H.fidx_first : H.flow4_first;
    }
    ++h->epoch;
    h->n_patched_buckets += dirty.size() - kImageHeaderBytes / 32;
    ++h->n_incremental_commits;
    rs.flow_log.clear();
    rs.l2_dirty = false;
    return 1;
}

}  // namespace

extern "C" {

int sfc_commit(sfc_handle* h) {
    if (!h) return -EINVAL;
    DeviceGuard g(h->device);
    int rc = incremental_commit(h, nullptr, false);
    if (rc != 0) return rc < 0 ? rc : 0;
    return full_commit(h);
}

int sfc_commit_stream(sfc_handle* h, void* cuda_stream) {
    if (!h) return -EINVAL;
    DeviceGuard g(h->device);
    int rc = incremental_commit(h, (cudaStream_t)cuda_stream, true);
    if (rc != 0) return rc < 0 ? rc : 0;
    return full_commit(h);
}

long sfc_stat(const sfc_handle* h, const char* name) {
    if (!h || !name) return -EINVAL;
    if (!strcmp(name, "full_commits")) return (long)h->n_full_commits;
    if (!strcmp(name, "incremental_commits")) return (long)h->n_incremental_commits;
    if (!strcmp(name, "patched_buckets") || !strcmp(name, "patched_sectors")) return (long)h->n_patched_buckets;
    if (!strcmp(name, "flows")) return (long)h->rs.flows.size();
    if (!strcmp(name, "flow4_first_bucket")) return (long)h->flow_first[0];
    if (!strcmp(name, "launches")) return (long)h->launches.load();
    return -ENOENT;
}

size_t sfc_table_size(const sfc_handle* h) { return h ? h->image_bytes : 0; }
uint64_t sfc_table_epoch(const sfc_handle* h) { return h ? h->epoch : 0; }

int sfc_table_export(const sfc_handle* hc, void* buf, size_t len) {
    sfc_handle* h = const_cast<sfc_handle*>(hc);
    if (!h || !buf) return -EINVAL;
    if (h->offline()) {
        if (h->host_image.empty()) return fail(h, -ENOENT, "sfc_table_export: nothing committed");
        if (len < h->host_image.size()) return fail(h, -ENOSPC, "sfc_table_export: buffer too small");
        memcpy(buf, h->host_image.data(), h->host_image.size());
        return 0;
    }
    if (!h->d_image) return fail(h, -ENOENT, "sfc_table_export: nothing committed");
    if (len < h->image_bytes) return fail(h, -ENOSPC, "sfc_table_export: buffer too small");
    DeviceGuard g(h->device);
    CUDA_TRY(h, cudaMemcpy(buf, h->d_image, h->image_bytes, cudaMemcpyDeviceToHost));
    return 0;
}

int sfc_table_import(sfc_handle* h, const void* buf, size_t len) {
    if (!h || !buf) return -EINVAL;
    if (const char* why = validate_image((const uint8_t*)buf, len)) return fail(h, -EINVAL, "sfc_table_import: %s", why);
    h->shadow_valid = false;  // an imported image is not described by this handle's rule store
    h->rs.other_dirty = true;
    if (h->offline()) {
        h->host_image.assign((const uint8_t*)buf, (const uint8_t*)buf + len);
        h->image_bytes = len;
        ImageHeader H;
        memcpy(&H, buf, sizeof(H));
        h->epoch = H.epoch;
        return 0;
    }
    h->host_image.clear();
    DeviceGuard g(h->device);
    return install_image_from_host(h, (const uint8_t*)buf, len);
}

int sfc_table_export_device(const sfc_handle* hc, void* dptr, size_t len, void* cuda_stream) {
    sfc_handle* h = const_cast<sfc_handle*>(hc);
    if (!h || !dptr) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device (there is no CPU data path)");
    if (!h->d_image) return fail(h, -ENOENT, "sfc_table_export_device: nothing committed");
    if (len < h->image_bytes) return fail(h, -ENOSPC, "sfc_table_export_device: buffer too small");
    DeviceGuard g(h->device);
    CUDA_TRY(h, cudaMemcpyAsync(dptr, h->d_image, h->image_bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)cuda_stream));
    return 0;
}

int sfc_table_import_device(sfc_handle* h, const void* dptr, size_t len, void* cuda_stream) {
    if (!h || !dptr || len < kImageHeaderBytes) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device (there is no CPU data path)");
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    // validate the header on the host before adopting the buffer
    std::vector<uint8_t> head(kImageHeaderBytes);
    CUDA_TRY(h, cudaMemcpyAsync(head.data(), dptr, kImageHeaderBytes, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(h, cudaStreamSynchronize(st));
    ImageHeader H;
    memcpy(&H, head.data(), sizeof(H));
    if (H.magic != kImageMagic || H.version != kImageVersion || H.total_bytes != len)
        return fail(h, -EINVAL, "sfc_table_import_device: bad magic, version or length");
    uint8_t* d = nullptr;
    CUDA_TRY(h, cudaMalloc(&d, len));
    cudaError_t e = cudaMemcpyAsync(d, dptr, len, cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
        cudaFree(d);
        return fail(h, -EIO, "sfc_table_import_device: %s", cudaGetErrorString(e));
    }
    h->shadow_valid = false;
    h->host_image.clear();
    h->rs.other_dirty = true;
    return publish_image(h, d, len, H);
}

// =================================================================== data plane ==========
int sfc_classify(sfc_handle* h, const void* hdr, const uint16_t* in_port, uint64_t* result, size_t n, uint32_t flags,
                 void* cuda_stream) {
    if (!h) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device (there is no CPU data path)");
    if (n == 0) return 0;
    if (!hdr || !in_port || !result) return fail(h, -EINVAL, "sfc_classify: null buffer");
    DeviceGuard g(h->device);
    return launch_classify(h, hdr, in_port, result, n, flags, (cudaStream_t)cuda_stream);
}

int sfc_classify_batches(sfc_handle* h, const sfc_batch* batches, uint32_t n_batches, uint32_t flags,
                         void* cuda_stream) {
    if (!h) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device (there is no CPU data path)");
    if (n_batches == 0) return 0;
    if (!batches) return fail(h, -EINVAL, "sfc_classify_batches: null batch list");
    // validate everything before the first launch: a bad descriptor must not leave half the list done
    for (uint32_t i = 0; i < n_batches; ++i) {
        const sfc_batch& b = batches[i];
        if (b.n == 0) continue;
        if (!b.hdr || !b.in_port || !b.result) return fail(h, -EINVAL, "sfc_classify_batches: null buffer in batch %u", i);
        if ((uintptr_t)b.hdr & 63) return fail(h, -EINVAL, "sfc_classify_batches: hdr of batch %u is not 64-byte aligned", i);
        if (b.n > (1u << 30)) return fail(h, -EINVAL, "sfc_classify_batches: batch %u has more than 2^30 packets", i);
    }
    DeviceGuard g(h->device);
    std::shared_lock<std::shared_mutex> lk(h->tmu);  // see sfc_handle::tmu
    if (!h->d_image) return fail(h, -EINVAL, "sfc_classify_batches: no committed table image (call sfc_commit)");
    const bool c = flags & SFC_F_COUNTERS, a = h->T.acl_nrules > 0;
    // Greedy split into launches: at most kMaxBatches descriptors, 2^31 tiles and 2^31 rows of span each
    // (TMA coordinates are signed 32-bit).
    constexpr uint64_t kMaxRows = (1ull << 31) - 2 * kTile;
    BatchList L;                       // on the stack: callers on different streams do not share it
    uintptr_t batch_hdr[kMaxBatches];  // header address of each entry until the lowest one is known
    uint32_t i = 0;
    while (i < n_batches) {
        L.count = 0;
        uintptr_t lo = 0, hi = 0;
        uint64_t tiles = 0;
        for (; i < n_batches && L.count < (uint32_t)kMaxBatches; ++i) {
            const sfc_batch& b = batches[i];
            if (b.n == 0) continue;
            const uintptr_t p = (uintptr_t)b.hdr, e = p + (uintptr_t)b.n * 64;
            const uintptr_t nlo = L.count ? std::min(lo, p) : p, nhi = L.count ? std::max(hi, e) : e;
            const uint64_t t = (b.n + kTile - 1) / kTile;
            if (L.count && ((nhi - nlo) / 64 > kMaxRows || tiles + t > (1ull << 31))) break;
            lo = nlo;
            hi = nhi;
            BatchDesc& d = L.b[L.count++];
            d.row0 = 0;  // filled in below, once the lowest address is known
            d.n = b.n;
            d.tile0 = (uint32_t)tiles;
            d.pad = 0;
            d.in_port = b.in_port;
            d.result = b.result;
            batch_hdr[L.count - 1] = p;
            tiles += t;
        }
        if (!L.count) break;  // only empty batches were left
        L.base = (const uint8_t*)lo;
        for (uint32_t k = 0; k < L.count; ++k) L.b[k].row0 = (uint32_t)((batch_hdr[k] - lo) / 64);
        const uint64_t rows = (hi - lo) / 64;
        int rc;
        const uint32_t nt = (uint32_t)tiles;
        cudaStream_t s = (cudaStream_t)cuda_stream;
        if (h->T.flow_format == kFlowFormatIndexed) {
            if (h->low_regs && !a && !c) rc = launch_multi<false, false, 5, 2>(h, 36, L, rows, nt, flags, s);
            else if (c && a) rc = launch_multi<true, true, 0, 2>(h, 35, L, rows, nt, flags, s);
            else if (c) rc = launch_multi<true, false, 0, 2>(h, 34, L, rows, nt, flags, s);
            else if (a) rc = launch_multi<false, true, 0, 2>(h, 33, L, rows, nt, flags, s);
            else rc = launch_multi<false, false, 0, 2>(h, 32, L, rows, nt, flags, s);
        } else if (h->low_regs && !a && !c) rc = launch_multi<false, false, 5>(h, 22, L, rows, nt, flags, s);
        else if (c && a) rc = launch_multi<true, true>(h, 19, L, rows, nt, flags, s);
        else if (c) rc = launch_multi<true, false>(h, 18, L, rows, nt, flags, s);
        else if (a) rc = launch_multi<false, true>(h, 17, L, rows, nt, flags, s);
        else rc = launch_multi<false, false>(h, 16, L, rows, nt, flags, s);
        if (rc) return rc;
    }
    return 0;
}

// The host path: chunks of a call go round a ring of staging buffers, one stream each — while chunk c is copied
// in, chunk c-1 is being classified and chunk c-2 copied out.  Submission and completion are separate so that a
// caller can queue the next batch before collecting the previous one: then the only thing PCIe ever waits for is
// the very first copy-in (a lone synchronous call also exposes its last kernel + copy-out).
namespace {
int host_retire(sfc_handle* h, int b) {  // wait for the chunk in ring slot b (its buffers may then be reused)
    if (!h->host_busy[b]) return 0;
    CUDA_TRY(h, cudaStreamSynchronize(h->hs[b]));
    h->host_busy[b] = false;
    if (h->host_timing) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[b][0], h->ev[b][1]);
        h->host_ms_acc += ms;
    }
    return 0;
}
}  // namespace

int sfc_classify_host_submit(sfc_handle* h, const void* hdr, const uint16_t* in_port, uint64_t* result, size_t n,
                             uint32_t flags, uint64_t* ticket) {
    if (!h) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device (there is no CPU data path)");
    if (ticket) *ticket = h->host_seq;
    if (n == 0) return 0;
    if (!hdr || !in_port || !result) return fail(h, -EINVAL, "sfc_classify_host: null buffer");
    DeviceGuard g(h->device);
    if (!h->hs[0]) {
        for (int i = 0; i < kHostBufs; ++i) {
            CUDA_TRY(h, cudaStreamCreateWithFlags(&h->hs[i], cudaStreamNonBlocking));
            CUDA_TRY(h, cudaEventCreate(&h->ev[i][0]));
            CUDA_TRY(h, cudaEventCreate(&h->ev[i][1]));
            CUDA_TRY(h, cudaMalloc(&h->d_hdr[i], kHostChunk * 64));
            CUDA_TRY(h, cudaMalloc(&h->d_port[i], kHostChunk * 2));
            CUDA_TRY(h, cudaMalloc(&h->d_res[i], kHostChunk * 8));
        }
    }
    const uint8_t* hp = (const uint8_t*)hdr;
    const size_t chunk = h->host_chunk;
    const bool timed = h->host_timing;  // two event records per chunk cost ~1.3 % of the call: opt-in
    for (size_t off = 0; off < n; off += chunk) {
        const size_t m = std::min(chunk, n - off);
        const int b = (int)(h->host_seq % (uint64_t)h->host_bufs);
        if (int rc = host_retire(h, b)) return rc;
        cudaStream_t st = h->hs[b];
        CUDA_TRY(h, cudaMemcpyAsync(h->d_hdr[b], hp + off * 64, m * 64, cudaMemcpyHostToDevice, st));
        CUDA_TRY(h, cudaMemcpyAsync(h->d_port[b], in_port + off, m * 2, cudaMemcpyHostToDevice, st));
        if (timed) CUDA_TRY(h, cudaEventRecord(h->ev[b][0], st));
        int rc = launch_classify(h, h->d_hdr[b], h->d_port[b], h->d_res[b], m, flags, st);
        if (rc) return rc;
        if (timed) CUDA_TRY(h, cudaEventRecord(h->ev[b][1], st));
        CUDA_TRY(h, cudaMemcpyAsync(result + off, h->d_res[b], m * 8, cudaMemcpyDeviceToHost, st));
        h->host_busy[b] = true;
        h->host_last[b] = h->host_seq++;
    }
    if (ticket) *ticket = h->host_seq;  // every chunk numbered below this belongs to this call or an earlier one
    return 0;
}

int sfc_classify_host_wait(sfc_handle* h, uint64_t ticket) {
    if (!h) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device (there is no CPU data path)");
    DeviceGuard g(h->device);
    // A slot whose latest chunk is newer than the ticket was reused since, and reuse waits for the older chunk:
    // only slots still holding a chunk of the awaited calls need waiting for.
    for (int b = 0; b < kHostBufs; ++b)
        if (h->host_busy[b] && h->host_last[b] < ticket)
            if (int rc = host_retire(h, b)) return rc;
    return 0;
}

int sfc_classify_host(sfc_handle* h, const void* hdr, const uint16_t* in_port, uint64_t* result, size_t n,
                      uint32_t flags) {
    if (!h) return -EINVAL;
    h->host_ms_acc = 0.f;
    uint64_t ticket = 0;
    if (int rc = sfc_classify_host_submit(h, hdr, in_port, result, n, flags, &ticket)) return rc;
    if (int rc = sfc_classify_host_wait(h, ticket)) return rc;
    if (n) h->last_kernel_ms = h->host_ms_acc;
    return 0;
}

int sfc_sync(sfc_handle* h, void* cuda_stream) {
    if (!h) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device");
    DeviceGuard g(h->device);
    CUDA_TRY(h, cudaStreamSynchronize((cudaStream_t)cuda_stream));
    return 0;
}

int sfc_counters_read(sfc_handle* h, uint64_t* out, size_t n) {
    if (!h || !out || n > SFC_CNT_TOTAL) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device");
    DeviceGuard g(h->device);
    CUDA_TRY(h, cudaDeviceSynchronize());
    CUDA_TRY(h, cudaMemcpy(out, h->d_counters, n * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return 0;
}

int sfc_counters_reset(sfc_handle* h) {
    if (!h) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device");
    DeviceGuard g(h->device);
    CUDA_TRY(h, cudaDeviceSynchronize());
    CUDA_TRY(h, cudaMemset(h->d_counters, 0, SFC_CNT_TOTAL * sizeof(uint64_t)));
    return 0;
}

void* sfc_counters_device_ptr(sfc_handle* h) { return h ? (void*)h->d_counters : nullptr; }
uint64_t sfc_launch_count(const sfc_handle* h) { return h ? h->launches.load() : 0; }
float sfc_last_kernel_ms(const sfc_handle* h) { return h ? h->last_kernel_ms : -1.f; }

// =================================================================== no-allocation call ==
// Scratch of sfc_classify_static: [0, 256) tile-ticket words of the classify kernel, then the table
// image.  One copy per GPU that loads this library, resident until the process exits.
static __device__ __align__(256) unsigned char g_static_area[SFC_STATIC_AREA_BYTES];

namespace {
constexpr size_t kStaticTicketBytes = 256;
constexpr int kStaticMaxDevices = 64;
struct StaticDev {
    bool ready = false;
    unsigned char* area = nullptr;  // device address of g_static_area ON THIS DEVICE
    int sm_count = 0;
    int ctas_per_sm[4] = {0, 0, 0, 0};  // [no ACL, ACL] x [sector flow tables, indexed flow table]
    EncodeTiledFn encode = nullptr;
    bool resident = false;          // an image was copied in by an earlier call
    ImageHeader header;
};
StaticDev g_static_dev[kStaticMaxDevices];

int static_fail(char* errbuf, size_t errlen, int code, const char* fmt, ...) {
    if (errbuf && errlen) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(errbuf, errlen, fmt, ap);
        va_end(ap);
    }
    return code;
}

// Once per device, with that device current.  In host code `g_static_area` is only a stand-in for
// the symbol, not a device address: the address a kernel may use comes from cudaGetSymbolAddress.
int static_setup(int device, StaticDev& d, char* errbuf, size_t errlen) {
    int major = 0;
    cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    if (e != cudaSuccess || major < 10)
        return static_fail(errbuf, errlen, -ENODEV, "device %d is not a compute-capability 10.x GPU", device);
    cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device);
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_area);
    if (e != cudaSuccess || !p)
        return static_fail(errbuf, errlen, -EIO, "cudaGetSymbolAddress(static area): %s", cudaGetErrorString(e));
    d.area = (unsigned char*)p;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn)
        return static_fail(errbuf, errlen, -ENOSYS, "driver lacks cuTensorMapEncodeTiled");
    d.encode = (EncodeTiledFn)fn;
    const void* kerns[4] = {(const void*)classify_kernel<kTile, kStages, false, false, false, 0, 1>,
                            (const void*)classify_kernel<kTileAcl, kStagesAcl, false, true, false, 0, 1>,
                            (const void*)classify_kernel<kTile, kStages, false, false, false, 0, 2>,
                            (const void*)classify_kernel<kTileAcl, kStagesAcl, false, true, false, 0, 2>};
    const int smem[2] = {SmemLayout<kTile, kStages, false>::kBytes, SmemLayout<kTileAcl, kStagesAcl, true>::kBytes};
    const int threads[2] = {kTile + 32, kTileAcl + 32};
    for (int v = 0; v < 4; ++v) {
        e = cudaFuncSetAttribute(kerns[v], cudaFuncAttributeMaxDynamicSharedMemorySize, smem[v & 1]);
        if (e == cudaSuccess)
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.ctas_per_sm[v], kerns[v], threads[v & 1], smem[v & 1]);
        if (e != cudaSuccess || d.ctas_per_sm[v] < 1)
            return static_fail(errbuf, errlen, -EIO, "classify kernel set-up: %s", cudaGetErrorString(e));
    }
    d.ready = true;
    return 0;
}
}  // namespace

extern "C" int sfc_classify_static(int device, const void* image_host, size_t image_bytes, const void* hdr,
                                   const uint16_t* in_port, uint64_t* result, size_t n, void* cuda_stream,
                                   char* errbuf, size_t errlen) {
    if (errbuf && errlen) errbuf[0] = 0;
    if (device < 0 || device >= kStaticMaxDevices)
        return static_fail(errbuf, errlen, -EINVAL, "device ordinal %d out of range", device);
    if (n > (1ull << 31)) return static_fail(errbuf, errlen, -EINVAL, "at most 2^31 packets per call");
    if (n && (!hdr || !in_port || !result)) return static_fail(errbuf, errlen, -EINVAL, "null packet buffer");
    if (((uintptr_t)hdr & 15) != 0) return static_fail(errbuf, errlen, -EINVAL, "hdr must be 16-byte aligned");
    // the size limit is checked before any CUDA call
    const size_t max_image = SFC_STATIC_AREA_BYTES - kStaticTicketBytes;
    if (image_host) {
        if (image_bytes > max_image)
            return static_fail(errbuf, errlen, -E2BIG,
                               "table image of %zu bytes does not fit the static device area: at most %zu bytes "
                               "(SFC_STATIC_AREA_BYTES %u minus %zu for kernel scratch); use sfc_classify with a handle",
                               image_bytes, max_image, (unsigned)SFC_STATIC_AREA_BYTES, kStaticTicketBytes);
        if (const char* why = validate_image((const uint8_t*)image_host, image_bytes))
            return static_fail(errbuf, errlen, -EINVAL, "%s", why);
    }
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device >= count)
        return static_fail(errbuf, errlen, -ENODEV, "no CUDA device %d (there is no CPU fallback)", device);
    DeviceGuard g(device);
    StaticDev& d = g_static_dev[device];
    if (!d.ready) {
        int rc = static_setup(device, d, errbuf, errlen);
        if (rc) return rc;
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    unsigned char* image_dev = d.area + kStaticTicketBytes;
    cudaError_t e;
    if (image_host) {
        e = cudaMemcpyAsync(image_dev, image_host, image_bytes, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return static_fail(errbuf, errlen, -EIO, "image copy: %s", cudaGetErrorString(e));
        memcpy(&d.header, image_host, sizeof(ImageHeader));
        d.resident = true;
    } else if (!d.resident) {
        return static_fail(errbuf, errlen, -ENOENT, "image_host is NULL but no earlier call left an image on device %d", device);
    }
    if (n == 0) return 0;
    // the array keeps the previous call's contents: the ticket words must start at zero
    e = cudaMemsetAsync(d.area, 0, kStaticTicketBytes, st);
    if (e != cudaSuccess) return static_fail(errbuf, errlen, -EIO, "scratch clear: %s", cudaGetErrorString(e));
    DevTables T{};
    fill_tables(T, d.header, image_dev, nullptr, -1);
    CUtensorMap map;
    cuuint64_t dims[2] = {16, (cuuint64_t)n};
    cuuint64_t strides[1] = {64};
    cuuint32_t box[2] = {16, (cuuint32_t)256};  // kTile and kTileAcl are multiples of 256 rows
    cuuint32_t estr[2] = {1, 1};
    CUresult r = d.encode(&map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(hdr), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return static_fail(errbuf, errlen, -EIO, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    const bool acl = T.acl_nrules > 0, indexed = T.flow_format == kFlowFormatIndexed;
    const int tile = acl ? kTileAcl : kTile;
    const uint64_t n_tiles = (n + tile - 1) / tile;
    const unsigned grid = (unsigned)std::min<uint64_t>(n_tiles, (uint64_t)d.sm_count * d.ctas_per_sm[(acl ? 1 : 0) + (indexed ? 2 : 0)]);
    uint32_t* ticket = (uint32_t*)d.area;  // {draw counter, finished producers}: cleared above, reset by the kernel too
#define SFC_STATIC_GO(A_, FMT_)                                                                                          \
    classify_kernel<(A_) ? kTileAcl : kTile, (A_) ? kStagesAcl : kStages, false, A_, false, 0, FMT_>                      \
        <<<grid, ((A_) ? kTileAcl : kTile) + 32, SmemLayout<(A_) ? kTileAcl : kTile, (A_) ? kStagesAcl : kStages, A_>::kBytes, st>>>( \
            map, in_port, result, (uint32_t)n, (uint32_t)n_tiles, ticket, 0u, T, NoBatches{})
    if (acl && indexed) SFC_STATIC_GO(true, 2);
    else if (acl) SFC_STATIC_GO(true, 1);
    else if (indexed) SFC_STATIC_GO(false, 2);
    else SFC_STATIC_GO(false, 1);
#undef SFC_STATIC_GO
    e = cudaGetLastError();
    if (e != cudaSuccess) return static_fail(errbuf, errlen, -EIO, "kernel launch: %s", cudaGetErrorString(e));
    return 0;
}

// =================================================================== generator ===========
int sfc_gen_packets(sfc_handle* h, const sfc_gen_spec* spec, uint64_t first, size_t n, void* hdr_dev,
                    uint16_t* in_port_dev, void* cuda_stream) {
    if (!h || !spec) return -EINVAL;
    if (h->offline()) return fail(h, -ENODEV, "offline handle: no device");
    if (n == 0) return 0;
    if (!hdr_dev || !in_port_dev) return fail(h, -EINVAL, "sfc_gen_packets: null buffer");
    DeviceGuard g(h->device);
    unsigned blocks = (unsigned)((n + 255) / 256);
    gen_kernel<<<blocks, 256, 0, (cudaStream_t)cuda_stream>>>(*spec, first, (uint64_t)n, (uint8_t*)hdr_dev, in_port_dev);
    CUDA_TRY(h, cudaGetLastError());
    ++h->launches;
    return 0;
}

void sfc_gen_packets_host(const sfc_gen_spec* spec, uint64_t first, size_t n, uint8_t* hdr, uint16_t* in_port) {
    for (size_t i = 0; i < n; ++i) gen_packet(*spec, first + i, hdr + i * 64, in_port[i]);
}

void sfc_gen_flows(const sfc_gen_spec* spec, uint32_t first, size_t n, sfc_flow_key* keys, sfc_flow_action* acts) {
    for (size_t i = 0; i < n; ++i) {
        GenFlow f;
        gen_flow(*spec, first + (uint32_t)i, f);
        sfc_flow_key& k = keys[i];
        memset(&k, 0, sizeof(k));
        for (int w = 0; w < (f.ip_ver == 6 ? 4 : 1); ++w) {
            put32(k.src + 4 * w, f.sip[w]);
            put32(k.dst + 4 * w, f.dip[w]);
        }
        k.sport = (uint16_t)f.sport;
        k.dport = (uint16_t)f.dport;
        k.vlan = (uint16_t)f.vlan;
        k.proto = (uint8_t)f.proto;
        k.ip_ver = (uint8_t)f.ip_ver;
        acts[i].action = (uint8_t)f.action;
        acts[i].reserved = 0;
        acts[i].out_port = (uint16_t)f.out_port;
    }
}

void sfc_gen_mac(uint32_t i, uint8_t mac[6]) { gen_mac(i, mac); }

void sfc_gen_prefix(const sfc_gen_spec* spec, int af, uint32_t j, uint8_t* prefix, int* len, uint16_t* out_port) {
    uint32_t w[4], port;
    int l;
    gen_prefix(*spec, af, j, w, l, port);
    for (int k = 0; k < (af == 6 ? 4 : 1); ++k) put32(prefix + 4 * k, w[k]);
    if (len) *len = l;
    if (out_port) *out_port = (uint16_t)port;
}

void sfc_gen_acl_rule(const sfc_gen_spec* spec, uint32_t j, sfc_acl_rule* r) {
    memset(r, 0, sizeof(*r));
    uint64_t a0 = draw(spec->seed_acl, 4ull * j), a1 = draw(spec->seed_acl, 4ull * j + 1);
    uint64_t a2 = draw(spec->seed_acl, 4ull * j + 2), a3 = draw(spec->seed_acl, 4ull * j + 3);
    bool v6 = spec->n_pfx6 && ((uint32_t)a0 % 10u) < 2u;
    r->ip_ver = v6 ? 6 : 4;
    // destination: a prefix of the LPM set, truncated to a coarse length, so traffic aimed at the
    // prefix set meets the rule
    uint32_t npfx = v6 ? spec->n_pfx6 : spec->n_pfx4;
    uint32_t w[4] = {(uint32_t)(a1 >> 32), (uint32_t)a1, (uint32_t)(a2 >> 32), (uint32_t)a2}, port;
    int plen = v6 ? 64 : 32;
    if (npfx) gen_prefix(*spec, v6 ? 6 : 4, (uint32_t)(a1 >> 32) % npfx, w, plen, port);
    const int cut4[4] = {8, 16, 24, 32}, cut6[4] = {32, 40, 48, 64};
    int dlen = std::min(plen, (v6 ? cut6 : cut4)[(a0 >> 8) & 3]);
    if (((a0 >> 12) & 7) == 0) dlen = 0;  // 1 in 8: any destination
    for (int k = 0; k < (v6 ? 4 : 1); ++k) {
        uint32_t m = mask32(dlen - 32 * k);
        put32(r->dst + 4 * k, w[k] & m);
        put32(r->dst_mask + 4 * k, m);
    }
    // source: mostly any, sometimes a short prefix
    uint32_t sl = (uint32_t)(a0 >> 16) & 15u;
    int slen = sl < 11 ? 0 : (sl < 14 ? 4 : 8);
    put32(r->src, (uint32_t)(a2 >> 32) & mask32(slen));
    put32(r->src_mask, mask32(slen));
    if ((a0 >> 20) & 1) {
        r->proto = ((a0 >> 21) & 1) ? 17 : 6;
        r->proto_mask = 0xFF;
    }
    uint32_t dp = (uint32_t)(a0 >> 24) & 15u;
    if (dp < 5) {  // exact destination port in a small well-known range
        r->dport = (uint16_t)((a3 >> 8) & 0x3FF);
        r->dport_mask = 0xFFFF;
    } else if (dp < 8) {  // a block of 4096 ports
        r->dport = (uint16_t)((a3 >> 8) & 0xF000);
        r->dport_mask = 0xF000;
    }
    if (((a0 >> 28) & 15u) == 0) {
        r->in_port = (uint16_t)(spec->port_base + ((uint32_t)(a3 >> 24) & 0xFFFF) % (spec->n_ports ? spec->n_ports : 1));
        r->in_port_mask = 0xFFFF;
    }
    r->verdict = ((uint32_t)a3 % 10u) == 0 ? SFC_ACL_DENY : SFC_ACL_PERMIT;
    // more specific rules outrank broader ones, with ties on purpose
    uint32_t spec_bits = (uint32_t)dlen + (r->dport_mask == 0xFFFF ? 16u : (r->dport_mask ? 4u : 0u)) + (uint32_t)slen;
    r->priority = 1u + spec_bits * 4u + (uint32_t)(a3 >> 32) % 4u;
}

}  // extern "C"
