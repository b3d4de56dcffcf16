/* A host process with no Python and no CUDA headers driving every GPU it is given through the C ABI alone — the
 * shape of the Go VSP of INTEGRATION.md, in C because no Go toolchain exists here.  Rules on rank 0, image broadcast
 * over NCCL, packets sharded, counters all-reduced; every GPU must give the answers rank 0 gives.
 * build: gcc -std=c11 -I include integration/c/group_demo.c -L dpu-operator_b200 -lb200sfc -Wl,-rpath,$PWD/dpu-operator_b200 -o /tmp/group_demo
 * run:   /tmp/group_demo <n_gpus>            (prints one line, exit code 0 = every check passed) */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "b200sfc.h"

#define CHECK(call)                                                                     \
    do {                                                                                \
        int rc_ = (call);                                                               \
        if (rc_ != 0) { fprintf(stderr, "%s -> %d (%s)\n", #call, rc_, g ? sfc_group_last_error(g) : ""); return 2; } \
    } while (0)

int main(int argc, char** argv) {
    int n_gpus = argc > 1 ? atoi(argv[1]) : 2;
    sfc_group* g = NULL;
    CHECK(sfc_group_create(n_gpus, NULL, &g));
    sfc_handle* h0 = sfc_group_handle(g, 0);

    sfc_gen_spec spec;
    memset(&spec, 0, sizeof spec);
    spec.seed_pkts = 0x5FC00002ull; spec.seed_rules = 0x5FC00001ull; spec.seed_lpm = 0x5FC00003ull; spec.seed_acl = 0x5FC00004ull;
    spec.n_flows = 100000; spec.n_ports = 8; spec.n_out_ports = 8; spec.n_macs = 8; spec.hit_permille = 800;
    sfc_flow_key* keys = calloc(spec.n_flows, sizeof *keys);
    sfc_flow_action* acts = calloc(spec.n_flows, sizeof *acts);
    sfc_gen_flows(&spec, 0, spec.n_flows, keys, acts);
    if (sfc_flow_add_bulk(h0, keys, acts, spec.n_flows, NULL)) { fprintf(stderr, "flow_add_bulk: %s\n", sfc_last_error(h0)); return 2; }
    if (sfc_commit(h0)) { fprintf(stderr, "commit: %s\n", sfc_last_error(h0)); return 2; }
    CHECK(sfc_group_broadcast_table(g));
    float bcast_ms = sfc_group_last_ms(g, 0);

    const size_t n = 1u << 20, shard = n / (size_t)n_gpus;
    uint8_t* hdr = malloc(n * 64);
    uint16_t* port = malloc(n * 2);
    uint64_t* res = calloc(n, 8);
    uint64_t* ref = calloc(n, 8);
    sfc_gen_packets_host(&spec, 0, n, hdr, port);
    if (sfc_classify_host(h0, hdr, port, ref, n, 0)) { fprintf(stderr, "classify on rank 0: %s\n", sfc_last_error(h0)); return 2; }
    for (int r = 0; r < n_gpus; ++r) {  /* each GPU its own contiguous range, counters on */
        size_t lo = (size_t)r * shard, m = r == n_gpus - 1 ? n - lo : shard;
        sfc_handle* h = sfc_group_handle(g, r);
        if (sfc_classify_host(h, hdr + lo * 64, port + lo, res + lo, m, SFC_F_COUNTERS)) { fprintf(stderr, "rank %d: %s\n", r, sfc_last_error(h)); return 2; }
    }
    if (memcmp(res, ref, n * 8) != 0) { fprintf(stderr, "sharded results differ from rank 0's\n"); return 3; }
    CHECK(sfc_group_reduce_counters(g));
    uint64_t hits = 0;
    for (int r = 0; r < n_gpus; ++r) {
        uint64_t c[16];
        if (sfc_counters_read(sfc_group_handle(g, r), c, 16)) return 2;
        uint64_t total = c[SFC_CNT_ACTION0] + c[SFC_CNT_ACTION0 + 1] + c[SFC_CNT_ACTION0 + 2];
        if (total != n) { fprintf(stderr, "rank %d counts %llu packets, expected %zu\n", r, (unsigned long long)total, n); return 4; }
        hits = c[SFC_CNT_STAGE0 + SFC_STAGE_FLOW];
    }
    printf("group_demo ok: %d GPUs, %zu packets, %llu flow hits, table %zu B, broadcast %.3f ms, all-reduce %.3f ms\n", n_gpus, n,
           (unsigned long long)hits, sfc_table_size(h0), bcast_ms, sfc_group_last_ms(g, 1));
    return sfc_group_destroy(g);
}
