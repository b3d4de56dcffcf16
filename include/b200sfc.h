/* b200sfc.h — C ABI of libb200sfc.so: batched flow classification + service-function-chain
 * steering on NVIDIA B200 (sm_100a).
 *
 * This is the drop-in boundary of the project.  The reference
 * (Ximinhan/dpu-operator, a Go control plane) never classifies a packet itself: its vendor
 * plugins *ask* a dataplane to do it, through the in-process Go interface `mrvldp`
 * (internal/daemon/vendor-specific-plugins/marvell/main.go:54-62) whose existing
 * implementations exec ovs-vsctl/ovs-ofctl (marvell/ovs-dp/ovsdp.go:40-162) or log and do
 * nothing (marvell/debug-dp/debugdp.go:1-50).  Every control-plane entry point below names the
 * reference call it replaces; the data-plane entry points (sfc_classify*) have no reference
 * counterpart because the reference's lookups happen in OVS / DPU hardware.
 *
 * Conventions
 *   - plain C, opaque handle, every call returns 0 on success or a negative errno-style code;
 *     sfc_last_error(h) returns a human-readable message for the last failure on that handle.
 *   - the caller owns every buffer it passes; the library owns all device memory it allocates.
 *   - control plane (everything except sfc_classify* / sfc_sync): single writer per handle,
 *     the library serialises nothing for you (the reference VSP mutates its stores without a
 *     lock as well, marvell/main.go:361-381).  Rule edits are staged on the host and become
 *     visible to the GPU only at sfc_commit(), which builds a new table image and swaps it in.
 *   - there is NO CPU fallback: if no CUDA device is present sfc_create fails with -ENODEV.
 *
 * Pinned per-packet semantics (the oracle in oracle/sfc_oracle.cpp is its executable definition):
 *   stage 0  parse the 64-byte header record (Ethernet, optional single 802.1Q tag, IPv4 with
 *            IHL honoured / IPv6 without extension headers, L4 ports for TCP/UDP/SCTP).
 *            Non-IP frames skip stages 1, 2 and 4.
 *   stage 1  ACL: ternary match, highest priority wins, ties -> lowest rule id.  deny => DROP,
 *            stop.  permit or no match => continue.
 *   stage 2  exact 5-tuple flow table (vlan, ip_ver, src, dst, proto, sport, dport). hit => stop.
 *   stage 3  SFC L2 steering: exact (in_port, dl_dst) first, then wildcard (in_port)
 *            (rule shapes of marvell/ovs-dp/ovsdp.go:113-142).  hit => FORWARD, stop.
 *   stage 4  LPM on the destination address (shape of ipv4_table / ipv6_table,
 *            cmd/intelvsp/fxp-net_linux-networking/fxp-net_linux-networking.p4info.txt:807,868).
 *   stage 5  miss => TO_CP (the default of l2_fwd_rx_table / l2_fwd_tx_table in the same p4info).
 */
#ifndef B200SFC_H
#define B200SFC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- data-plane records ---- */

#define SFC_HDR_BYTES 64u /* one header record: the first 64 bytes of the L2 frame, zero padded */

/* result word: action:8 | next_nf:8 | out_port:16 | rule_id:32  (most significant first) */
#define SFC_ACT_TO_CP 0u   /* miss: punt to the control plane */
#define SFC_ACT_FORWARD 1u /* send to out_port */
#define SFC_ACT_DROP 2u
#define SFC_NF_NONE 0xFFu /* next_nf when out_port belongs to no network function */

/* rule_id = stage:4 | index:28 ; index is the id the matching sfc_*_add call returned */
#define SFC_STAGE_MISS 0u
#define SFC_STAGE_ACL 1u
#define SFC_STAGE_FLOW 2u
#define SFC_STAGE_L2_EXACT 3u
#define SFC_STAGE_L2_WILD 4u
#define SFC_STAGE_LPM4 5u
#define SFC_STAGE_LPM6 6u

#define SFC_RESULT(action, next_nf, out_port, stage, index)                                   \
    (((uint64_t)(action) << 56) | ((uint64_t)(next_nf) << 48) | ((uint64_t)(out_port) << 32) | \
     ((uint64_t)(stage) << 28) | (uint64_t)(index))
#define SFC_RESULT_MISS SFC_RESULT(SFC_ACT_TO_CP, SFC_NF_NONE, 0, SFC_STAGE_MISS, 0)
#define SFC_RESULT_ACTION(r) ((unsigned)((r) >> 56) & 0xFFu)
#define SFC_RESULT_NEXT_NF(r) ((unsigned)((r) >> 48) & 0xFFu)
#define SFC_RESULT_OUT_PORT(r) ((unsigned)((r) >> 32) & 0xFFFFu)
#define SFC_RESULT_STAGE(r) ((unsigned)((r) >> 28) & 0xFu)
#define SFC_RESULT_INDEX(r) ((unsigned)(r) & 0x0FFFFFFFu)

/* limits */
#define SFC_MAX_PORTS 4096u          /* port ids are 0 .. SFC_MAX_PORTS-1 */
#define SFC_MAX_FLOWS (1u << 26)     /* flow ids are 26 bits wide in a table slot */
#define SFC_MAX_ACL_RULES 4096u
#define SFC_MAX_NFS 254u

/* 5-tuple flow key.  Addresses are in network byte order; an IPv4 address occupies bytes 0..3
 * and bytes 4..15 must be zero.  Ports, vlan are host-order integers. */
typedef struct sfc_flow_key {
    uint8_t src[16];
    uint8_t dst[16];
    uint16_t sport;
    uint16_t dport;
    uint16_t vlan;  /* 0 = untagged, else the 12-bit VID */
    uint8_t proto;  /* IPv4 protocol / IPv6 next header */
    uint8_t ip_ver; /* 4 or 6 */
} sfc_flow_key;     /* 40 bytes */

typedef struct sfc_flow_action {
    uint8_t action; /* SFC_ACT_* */
    uint8_t reserved;
    uint16_t out_port; /* used when action == SFC_ACT_FORWARD, else must be 0 */
} sfc_flow_action;     /* 4 bytes */

/* Ternary ACL rule: a packet field f matches when (f & mask) == (value & mask).  ip_ver 0 = any
 * version (then the address masks must be zero), 4 or 6 = only that version.  Priority semantics
 * are P4Runtime's (cmd/intelvsp/p4rt-ctl:455-534): larger number wins; ties -> lowest rule id. */
#define SFC_ACL_PERMIT 0u
#define SFC_ACL_DENY 1u
typedef struct sfc_acl_rule {
    uint8_t src[16], src_mask[16];
    uint8_t dst[16], dst_mask[16];
    uint16_t in_port, in_port_mask;
    uint16_t vlan, vlan_mask;
    uint16_t sport, sport_mask;
    uint16_t dport, dport_mask;
    uint8_t proto, proto_mask;
    uint8_t ip_ver;  /* 0, 4 or 6 */
    uint8_t verdict; /* SFC_ACL_PERMIT / SFC_ACL_DENY */
    uint32_t priority;
} sfc_acl_rule; /* 88 bytes */

/* counters block returned by sfc_counters_read (all uint64_t) */
#define SFC_CNT_ACTION0 0u   /* [0..2]  packets per action (TO_CP, FORWARD, DROP) */
#define SFC_CNT_STAGE0 4u    /* [4..10] packets decided per stage (SFC_STAGE_*) */
#define SFC_CNT_RULES0 16u   /* [16 .. 16+SFC_MAX_ACL_RULES) per ACL rule id, then            */
#define SFC_CNT_L2_0 (SFC_CNT_RULES0 + SFC_MAX_ACL_RULES) /* per L2 rule id (mod SFC_CNT_L2_N) */
#define SFC_CNT_L2_N 16384u
#define SFC_CNT_TOTAL (SFC_CNT_L2_0 + SFC_CNT_L2_N)

typedef struct sfc_handle sfc_handle;

/* ---------------------------------------------------------------- life cycle ------------ */

/* Create a dataplane instance on CUDA device `device_ordinal`.
 * Replaces: mrvldp.InitDataPlane(bridgeName) — marvell/main.go:58, ovs-dp/ovsdp.go:79-102
 * (there: `ovs-vsctl add-br`; here: bind a device, create an empty table image). */
int sfc_create(int device_ordinal, sfc_handle** out);
/* device_ordinal -1 creates an OFFLINE handle: rules can be staged, sfc_commit builds the image
 * in host memory and sfc_table_export returns it (so a control-plane host without a GPU can
 * prepare an image for the GPU ranks), but every data-plane call fails with -ENODEV. */

/* Free every host and device resource.
 * Replaces: mrvldp.DeleteDataplane(bridgeName) — marvell/main.go:60, ovs-dp/ovsdp.go:66-69. */
int sfc_destroy(sfc_handle* h);

/* Message for the last error on this handle (never NULL; valid until the next failing call).
 * Replaces: the Go `error` value every mrvldp method returns (marvell/main.go:54-62). */
const char* sfc_last_error(const sfc_handle* h);

/* Library version string "b200sfc <major>.<minor>" and the image format number. */
const char* sfc_version(void);

/* Tuning knobs.  No reference counterpart (the reference has no table).  Table-shape options take effect at the
 * next sfc_commit, which then rebuilds the image:
 *   "flow_format"     0 by size (default; picks 1), 1 one-sector cuckoo tables per address family, 2 fingerprint
 *                     index + dense records (half the memory, slower: profiles/r02_experiments.md)
 *   "flow_load_pct"   0 = the format's default (40 / 65), else 5..90: load factor the flow tables are built at
 *   "flow_hashes"     2 (default) or 3 bucket choices per flow (format 1)
 *   "flow_slots"      0 / 1 one slot per bucket (default), 2 dense two-slot buckets (format 1)
 * Launch-policy options take effect at once:
 *   "flow_probe_mode" -1 by table size (default); bit 0 = issue both probes at once, bit 1 = probes may stay in L1
 *   "low_regs"        -1 by table size (default), 0 / 1: run the 40-register kernels (five CTAs per SM)
 * sfc_classify_host pipeline:
 *   "host_chunk"      packets per chunk (1024 .. 262144, default 262144)
 *   "host_bufs"       chunks in flight (2 .. 4, default 2)
 *   "host_timing"     1 = record kernel time per chunk for sfc_last_kernel_ms (costs ~1 %)
 * Unknown names return -ENOENT.  (The experiment build, `make tune`, adds tune_* options; they are not part of this ABI.) */
int sfc_set_option(sfc_handle* h, const char* name, long value);

/* ---------------------------------------------------------------- port registry --------- */

/* Register a port by name and get its numeric id (the `in_port` value packets carry).
 * Adding an existing name returns its id (ovs-vsctl --may-exist semantics).
 * Replaces: mrvldp.AddPortToDataPlane(bridge, portName, vfPCI, isDPDK) —
 * marvell/main.go:55, ovs-dp/ovsdp.go:40-55. */
int sfc_port_add(sfc_handle* h, const char* name, uint16_t* port_id);

/* Remove a port; -ENOENT if unknown.  The L2 steering rules that name the port (as in_port or out_port) are
 * removed with it: with OVS a flow that names a deleted port stops matching, and here the id is handed out again
 * by the next sfc_port_add, so a rule left behind would steer the NEW port's traffic.  Takes effect at sfc_commit.
 * Replaces: mrvldp.DeletePortFromDataPlane — marvell/main.go:56, ovs-dp/ovsdp.go:58-62. */
int sfc_port_del(sfc_handle* h, const char* name);

/* Look a port up by name; -ENOENT if unknown. */
int sfc_port_lookup(const sfc_handle* h, const char* name, uint16_t* port_id);

/* Write the registered port names, one per line in id order, NUL-terminated, into buf.
 * Returns the number of bytes needed (including NUL); call with len 0 to size the buffer.
 * Replaces: mrvldp.ReadAllPortFromDataPlane — marvell/main.go:59, ovs-dp/ovsdp.go:104-112
 * (`ovs-vsctl list-ports`). */
long sfc_port_list(const sfc_handle* h, char* buf, size_t len);

/* ---------------------------------------------------------------- SFC L2 steering ------- */

/* Add one steering rule.  Shapes follow ovs-dp/ovsdp.go:113-142 exactly:
 *   dst_mac == NULL                     -> "priority=10,in_port=S,actions=output:D"      (R1)
 *   dst_mac != NULL, in_port != out_port-> "in_port=S,dl_dst=M,actions=output:D" (prio 32768, R2)
 *   dst_mac != NULL, in_port == out_port-> "priority=100,in_port=S,dl_dst=M,actions=in_port" (R3)
 * Re-adding an identical match replaces the action (ovs-ofctl add-flow semantics) and keeps
 * the rule id.  *rule_id (may be NULL) receives the id reported in result words.
 * Replaces: mrvldp.AddFlowRuleToDataPlane(bridge, inPort, outPort, dstMac) — marvell/main.go:61. */
int sfc_l2_rule_add(sfc_handle* h, uint16_t in_port, uint16_t out_port, const uint8_t* dst_mac,
                    uint32_t* rule_id);

/* Delete by match, as `ovs-ofctl del-flows` does (ovs-dp/ovsdp.go:143-162): with dst_mac the
 * rules matching (in_port, dl_dst); with dst_mac == NULL *every* rule whose in_port is S —
 * wildcard and MAC-qualified alike (a loose OpenFlow delete).  Deleting nothing is not an error.
 * Replaces: mrvldp.DeleteFlowRuleFromDataPlane — marvell/main.go:62. */
int sfc_l2_rule_del(sfc_handle* h, uint16_t in_port, const uint8_t* dst_mac);

/* Declare the chain: NF i owns ports nf_in_ports[i] and nf_out_ports[i]; result.next_nf is the
 * index of the NF owning result.out_port, or SFC_NF_NONE.
 * Reference anchor: ServiceFunctionChainSpec.NetworkFunctions (api/v1/servicefunctionchain_types.go:27-34),
 * each NF having exactly two DPU ports (internal/daemon/sfc-reconciler/sfc.go:39,55). */
int sfc_chain_set(sfc_handle* h, const uint16_t* nf_in_ports, const uint16_t* nf_out_ports,
                  size_t n_nfs);

/* ---------------------------------------------------------------- exact 5-tuple flows --- */

/* No reference counterpart; key shape after ipsec_spd / ipv4_tunnel_term
 * exact tables of the p4info (fxp-net_linux-networking.p4info.txt:507,1261). */
int sfc_flow_add(sfc_handle* h, const sfc_flow_key* key, const sfc_flow_action* act,
                 uint32_t* flow_id);
/* n keys at once; ids are *first_id .. *first_id+n-1 in array order.  Duplicate keys inside the
 * batch or against the table are rejected with -EEXIST and nothing is added. */
int sfc_flow_add_bulk(sfc_handle* h, const sfc_flow_key* keys, const sfc_flow_action* acts,
                      size_t n, uint32_t* first_id);
int sfc_flow_del(sfc_handle* h, const sfc_flow_key* key);

/* ---------------------------------------------------------------- LPM ------------------- */

/* af = 4 or 6; prefix = 4 or 16 bytes, network order, bits beyond `len` ignored.
 * Shape: ipv4_table (exact 32 + LPM 32) / ipv6_table (LPM 128), p4info.txt:807,868, with the
 * action collapsed to "forward to the port of the next hop" (ipv4_set_nexthop_id :2294). */
int sfc_lpm_add(sfc_handle* h, int af, const uint8_t* prefix, int len, uint16_t out_port,
                uint32_t* prefix_id);
int sfc_lpm_del(sfc_handle* h, int af, const uint8_t* prefix, int len);

/* ---------------------------------------------------------------- ACL ------------------- */

/* Ternary + priority, after the p4info's ternary tables (tx_source_port :1062,
 * source_port_to_bridge_map :981) and p4rt-ctl's priority rule (cmd/intelvsp/p4rt-ctl:455-534). */
int sfc_acl_add(sfc_handle* h, const sfc_acl_rule* rule, uint32_t* rule_id);
int sfc_acl_del(sfc_handle* h, uint32_t rule_id);

/* Remove every rule, flow, prefix and chain declaration (ports stay).  Ids restart at 0.
 * Reference anchor: the Intel VSP flushing its rules on exit (vendored ipuplugin/ipuplugin.go:230). */
int sfc_clear(sfc_handle* h);

/* ---------------------------------------------------------------- commit / image -------- */

/* Make the staged rule set the one sfc_classify uses.  No reference counterpart: OVS applies each `add-flow`
 * immediately; here edits are batched into epochs.  What a commit costs depends on what changed since the last one:
 *   - flows added / deleted, L2 steering rules added / deleted (what the reference really churns: three rules per
 *     bridge port, marvell/main.go:383-403, 441-460): the sectors of the committed image that changed are patched
 *     in place (flows) or written into the idle one of two copies of the L2 tables, which then becomes the active
 *     one.  Tens of microseconds; no rebuild, no allocation once the staging buffers exist.
 *   - anything else (LPM, ACL, the chain, table-shape options), or a table that has to grow: the image is rebuilt
 *     on the host and uploaded next to the old one.
 * Classify calls may run on other threads and streams while a commit is in progress.  A launch sees exactly one
 * table descriptor: the new image (or L2 copy) is published only after it has landed on the device, and the old
 * image is released only after every launch that could still read it has finished (the committing thread waits
 * for the device; data-plane threads are not held up).  Flow patches are the exception: they change the live
 * table sector by sector, so a launch running beside the commit may see a flow either present or absent per
 * packet (each packet's answer is that of the rule set before or after the commit, never anything else), and a
 * flow that the insertion moved between its two buckets can be missed by such a launch; order commits against
 * launches with sfc_commit_stream or a stream wait if that matters.
 * If a commit fails (for example no device memory for the second image) the committed image, the staged rules and
 * the edit log are unchanged, and the next commit tries again. */
int sfc_commit(sfc_handle* h);

/* Same, ordered on a stream instead of synchronising the device: when only flows were added or
 * deleted since the last commit (and the table has room), the edits are applied to a host shadow of
 * the cuckoo table and ONLY the touched buckets are shipped and written in place by a small kernel
 * on `cuda_stream` — O(edits), no rebuild, no device-wide sync; classify launches queued behind it on
 * the same stream see the new flows, launches before it the old ones.  Anything else (L2 rules,
 * prefixes, ACL, chain, options, a table that must grow) falls back to the full rebuild of sfc_commit.
 * sfc_commit itself takes the same incremental path when it can, then waits for it.
 * Reference anchor: per-pod rule churn — CreateBridgePort / DeleteBridgePort adding and removing
 * three flows each (marvell/main.go:383-403, 441-460) — which OVS applies one `add-flow` at a time. */
int sfc_commit_stream(sfc_handle* h, void* cuda_stream);

/* Counters of the library itself (not of traffic): "full_commits", "incremental_commits",
 * "patched_sectors" (32-byte sectors shipped by incremental commits; "patched_buckets" is the same number under its
 * older name), "flows", "flow4_first_bucket", "launches".  Negative errno for unknown names.
 * Reference anchor: the health/stat surface of DeviceService.GetDevices (marvell/main.go:219-236). */
long sfc_stat(const sfc_handle* h, const char* name);

/* Size in bytes of the current committed image, and its epoch (number of commits so far). */
size_t sfc_table_size(const sfc_handle* h);
uint64_t sfc_table_epoch(const sfc_handle* h);

/* Copy the committed image to / from caller memory.  The *_device variants take device
 * pointers on the handle's device so an NCCL broadcast can move the image rank to rank without
 * a host bounce.  An imported image replaces the committed one (staged rules are untouched).
 * sfc_table_import checks everything the kernel follows — section bounds and alignment, every index of the
 * LPM tries, the ACL forest and the flow index, the fill of the L2 table, the action and stage of every stored
 * result — and refuses a damaged image with -EINVAL.  sfc_table_import_device checks the header only (magic,
 * version, length): it is meant for images another handle of this library exported (rank-to-rank broadcast);
 * bring an image of unknown origin in through sfc_table_import.
 * No reference counterpart (single-DPU control plane). */
int sfc_table_export(const sfc_handle* h, void* buf, size_t len);
int sfc_table_import(sfc_handle* h, const void* buf, size_t len);
int sfc_table_export_device(const sfc_handle* h, void* dptr, size_t len, void* cuda_stream);
int sfc_table_import_device(sfc_handle* h, const void* dptr, size_t len, void* cuda_stream);

/* ---------------------------------------------------------------- data plane ------------ */

#define SFC_F_COUNTERS 1u /* also accumulate the SFC_CNT_* counters */
/* SFC_F_OVERLAP: the caller promises that this launch and the classify launch submitted just
 * before it on the same stream touch different result buffers and do not depend on one another.
 * The library then lets the new launch start while the previous one is still draining
 * (programmatic dependent launch); stream order towards every later operation is kept.
 * Meant for a STREAM of launches: each one uses a small grid so that several are resident at
 * once; a single isolated launch is faster without the flag. */
#define SFC_F_OVERLAP 2u

/* Classify n packets.  hdr (n x 64 B, 16-byte aligned), in_port (n x u16) and result (n x u64)
 * are DEVICE pointers on the handle's device.  Asynchronous on `cuda_stream` (a cudaStream_t,
 * NULL = legacy default stream).  One caller thread per stream; threads on different streams may
 * share the handle (up to 256 classify launches in flight per handle).
 * The call allocates nothing and synchronises nothing, so — after one launch outside capture, which
 * sets the kernel's attributes — it and sfc_classify_batches may be captured into a CUDA graph and
 * replayed (tests/test_gpu_parity.py).
 * No reference counterpart — this is the lookup the reference delegates to OVS / FXP hardware. */
int sfc_classify(sfc_handle* h, const void* hdr, const uint16_t* in_port, uint64_t* result,
                 size_t n, uint32_t flags, void* cuda_stream);

/* Many independent batches in ONE launch (BASELINE.json configs[2]: 4096-packet batches).  A NIC
 * poll loop hands over bursts far smaller than what keeps 148 SMs busy, and one launch per burst
 * is bound by launch latency; this call takes the bursts as they are — separate buffers, ragged
 * sizes — and classifies them as if they were one batch.  `batches` is a HOST array (read before
 * the call returns, nothing is retained); the buffers it names are DEVICE pointers, each hdr
 * 64-byte aligned.  Results are exactly those of one sfc_classify per batch.  Lists longer than
 * 512 entries are split into several launches.  Nothing is allocated or copied: the list travels
 * as a kernel parameter.  No reference counterpart (see sfc_classify). */
typedef struct sfc_batch {
    const void* hdr;         /* n x 64 B */
    const uint16_t* in_port; /* n x u16 */
    uint64_t* result;        /* n x u64 */
    uint32_t n;              /* packets; 0 = skip */
    uint32_t reserved;
} sfc_batch;
int sfc_classify_batches(sfc_handle* h, const sfc_batch* batches, uint32_t n_batches, uint32_t flags,
                         void* cuda_stream);

/* Same, with HOST buffers (pageable or pinned): the library stages them through device memory it
 * owns, runs the kernel and copies the results back before returning.  This is the call a host
 * dataplane shim (cgo) makes per batch. */
int sfc_classify_host(sfc_handle* h, const void* hdr, const uint16_t* in_port, uint64_t* result,
                      size_t n, uint32_t flags);

/* The same in two steps, for a loop that queues batch k+1 before it collects batch k (the copy engines then
 * never wait for the host).  submit returns once the work is queued and writes a ticket; the three buffers
 * must stay valid and untouched until sfc_classify_host_wait(h, ticket) returns, which it does once that
 * call and every earlier one have completed (ticket ~0 = everything).  One caller thread per handle. */
int sfc_classify_host_submit(sfc_handle* h, const void* hdr, const uint16_t* in_port,
                             uint64_t* result, size_t n, uint32_t flags, uint64_t* ticket);
int sfc_classify_host_wait(sfc_handle* h, uint64_t ticket);

/* Block until everything submitted on `cuda_stream` has finished. */
int sfc_sync(sfc_handle* h, void* cuda_stream);

/* Copy the first n counters (n <= SFC_CNT_TOTAL) to `out` (host), after syncing the device. */
int sfc_counters_read(sfc_handle* h, uint64_t* out, size_t n);
int sfc_counters_reset(sfc_handle* h);
/* Device address of the SFC_CNT_TOTAL x u64 counter array (for an NCCL all-reduce in place). */
void* sfc_counters_device_ptr(sfc_handle* h);

/* Number of kernel launches this handle has made (bench.py's gpu_launches claim). */
uint64_t sfc_launch_count(const sfc_handle* h);

/* Device time of the classify kernels launched by the last sfc_classify_host call, in ms
 * (CUDA events on the library's streams).  Measured only while sfc_set_option(h, "host_timing", 1)
 * is in force (two event records per chunk are not free); 0 otherwise, < 0 if no call was made. */
float sfc_last_kernel_ms(const sfc_handle* h);

/* ---------------------------------------------------------------- several GPUs ----------
 * One process, several B200s — the shape of a Go VSP.  The path shards by packets:
 * the table image is replicated, every GPU classifies its own batches with the calls above on
 * its own handle, and there is no collective on the per-batch path.  The two collectives that
 * exist run over NCCL (NVLink / NVSwitch), opened with dlopen("libnccl.so.2") at creation:
 *   sfc_group_broadcast_table  rank 0's committed image -> every other GPU (ncclBroadcast);
 *   sfc_group_reduce_counters  sum of the SFC_CNT_TOTAL counters over the GPUs, in place on each
 *                              (ncclAllReduce): every array then holds the total, so reset before
 *                              counting again.
 * Rules are staged and committed on rank 0's handle only (sfc_group_handle(g, 0)); the other
 * handles receive images.  device_ordinals = NULL means GPUs 0 .. n_gpus-1.  Errors: negative
 * errno, text from sfc_group_last_error; -ENODEV without a GPU, -ENOSYS if NCCL cannot be opened.
 * sfc_group_last_ms: device time of the last broadcast (which = 0) or all-reduce (which = 1).
 * No reference counterpart (the reference programs one DPU per VSP process). */
typedef struct sfc_group sfc_group;
int sfc_group_create(int n_gpus, const int* device_ordinals, sfc_group** out);
int sfc_group_destroy(sfc_group* g);
int sfc_group_size(const sfc_group* g);
sfc_handle* sfc_group_handle(sfc_group* g, int rank);
int sfc_group_broadcast_table(sfc_group* g);
int sfc_group_reduce_counters(sfc_group* g);
float sfc_group_last_ms(const sfc_group* g, int which);
const char* sfc_group_last_error(const sfc_group* g);

/* ---------------------------------------------------------------- no-allocation call ---- */

/* Classify without ever allocating.  This entry point has no counterpart in the reference.
 *
 * Who needs it: callers that are not allowed to obtain device memory at the moment they have
 * packets to classify — a dataplane thread that runs under a fixed memory budget handed out at
 * start-up, a shim loaded into a process whose allocator is owned by another runtime, a polling
 * loop that must not take the driver's allocation locks (cudaMalloc/cudaFree serialise the whole
 * device).  sfc_create/sfc_commit call cudaMalloc for the table image and the kernel's scratch;
 * this call does neither.  It takes no handle: the caller passes the table image (as built by an
 * OFFLINE handle's sfc_commit + sfc_table_export, or exported from any handle) in HOST memory
 * together with the packet buffers, and the library keeps what it needs on the GPU in a fixed
 * array that is part of the library image itself:
 *
 *   SFC_STATIC_AREA_BYTES of device memory per GPU, held on every GPU that loads the library until
 *   the process exits.  The CUDA runtime sets the array up when the library's code is first loaded
 *   onto a device — with lazy module loading that is the first call here, or any earlier kernel
 *   launch of this library, on that device — so that one call can fail for lack of memory; no later
 *   call can.  The first 256 bytes are the classify kernel's tile-ticket
 *   scratch; the rest holds the table image, so an image may be at most
 *   SFC_STATIC_AREA_BYTES - 256 bytes (a steering set plus tens of thousands of flows).  A larger
 *   image is refused with -E2BIG and a message naming both sizes; nothing is truncated.
 *
 * Behaviour
 *   - hdr / in_port / result are DEVICE pointers as for sfc_classify; image_host is a HOST pointer.
 *     Pass image_host == NULL to classify against the image the previous call left on that device.
 *   - asynchronous on cuda_stream; results are bit-identical to sfc_classify with the same image.
 *     (An image_host in pageable memory makes the copy wait for earlier work on the stream and return
 *     only when the bytes have left the buffer; pass pinned memory, or NULL, to stay asynchronous.)
 *   - the FIRST call on each device also does one-time set-up (looks up the array's device
 *     address, raises the kernel's shared-memory limit).  Make that first call before starting any
 *     stream capture on the device; later calls only copy, clear and launch.
 *   - every call on a device shares the one array and it keeps what the last call left there.  The
 *     call clears the ticket words itself (they must start at zero) and overwrites the image when
 *     one is passed; it does not and cannot protect the array from a second call running at the
 *     same time.  Two calls whose GPU work overlaps — different streams, or threads — race on the
 *     tickets and on the image: run them on one stream, or order them with events.
 *   - counters (SFC_F_COUNTERS) are not available here: they live in handle-owned memory.
 *   - on failure a negative errno is returned and, if errbuf != NULL, a NUL-terminated message of
 *     at most errlen bytes is written to it.  Never calls cudaMalloc, cudaFree, or anything that
 *     allocates on its behalf. */
#define SFC_STATIC_AREA_BYTES (4u << 20)
int sfc_classify_static(int device, const void* image_host, size_t image_bytes, const void* hdr,
                        const uint16_t* in_port, uint64_t* result, size_t n, void* cuda_stream,
                        char* errbuf, size_t errlen);

/* ---------------------------------------------------------------- synthetic traffic ----- */

/* Fill device buffers with the deterministic synthetic workload of SURVEY.md §8(d) (counter-based
 * splitmix64; the same workload is written a second time, independently, in oracle/workload.cpp, and
 * tests/test_oracle.py holds the two to identical bytes).
 * `spec` is an sfc_gen_spec; packets [first, first+n) of the stream are produced. */
typedef struct sfc_gen_spec {
    uint64_t seed_pkts;      /* packet stream seed                          */
    uint64_t seed_rules;     /* seed the flow rule set is drawn from        */
    uint64_t seed_lpm;       /* seed the prefix set is drawn from           */
    uint64_t seed_acl;       /* seed the ACL rule set is drawn from         */
    uint32_t n_flows;        /* flows 0..n_flows-1 exist (flow i = sfc_gen_flow(spec, i)) */
    uint32_t n_ports;        /* in_port drawn from port_base .. port_base+n_ports-1 */
    uint32_t port_base;
    uint32_t n_out_ports;    /* flow / next-hop out_port drawn from 0..n_out_ports-1 */
    uint32_t hit_permille;   /* share of IP packets that replay a flow of the set */
    uint32_t nonip_permille; /* share of frames that are not IP (ARP-like) */
    uint32_t v6_permille;    /* share of flows / random IP packets that are IPv6 */
    uint32_t vlan_permille;  /* share of flows / frames carrying an 802.1Q tag */
    uint32_t zipf;           /* 0 uniform over flows, 1 log-uniform octaves (Zipf, s ~ 1) */
    uint32_t n_macs;         /* dst MAC = sfc_gen_mac(k), k uniform in 0..n_macs-1 */
    uint32_t n_pfx4;         /* prefixes 0..n-1 exist (sfc_gen_prefix(spec, 4, j)) */
    uint32_t n_pfx6;
    uint32_t lpm_permille;   /* share of non-flow IP packets whose dst is drawn from a prefix */
    uint32_t n_acl;          /* ACL rules 0..n_acl-1 exist (sfc_gen_acl_rule(spec, j)) */
} sfc_gen_spec;

/* Fill DEVICE buffers with packets [first, first+n) of the stream (a kernel; async on stream). */
int sfc_gen_packets(sfc_handle* h, const sfc_gen_spec* spec, uint64_t first, size_t n,
                    void* hdr_dev, uint16_t* in_port_dev, void* cuda_stream);
/* Host-side pure functions of the same generator (no GPU needed; same source as the kernel). */
void sfc_gen_packets_host(const sfc_gen_spec* spec, uint64_t first, size_t n, uint8_t* hdr,
                          uint16_t* in_port);
void sfc_gen_flows(const sfc_gen_spec* spec, uint32_t first, size_t n, sfc_flow_key* keys,
                   sfc_flow_action* acts);
void sfc_gen_mac(uint32_t i, uint8_t mac[6]);
/* prefix j of family af (4|6): writes 4 or 16 prefix bytes, its length and next-hop port */
void sfc_gen_prefix(const sfc_gen_spec* spec, int af, uint32_t j, uint8_t* prefix, int* len,
                    uint16_t* out_port);
void sfc_gen_acl_rule(const sfc_gen_spec* spec, uint32_t j, sfc_acl_rule* rule);

#ifdef __cplusplus
}
#endif
#endif /* B200SFC_H */
