"""Measure BASELINE.json configs[2..4] on one B200 (parity for them lives in tests/; this tool only
times).  Writes one JSON object per line.  usage: python tools/bench_configs.py [cfg3 cfg4 cfg5] """
import importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from common import Scenario, config3_spec, config4_spec, install_flows, install_lpm_acl, sfc

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
OVL = sfc.consts.F_OVERLAP
st = torch.cuda.current_stream().cuda_stream


def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1e-3


def report(name, mode, pkts, secs, extra=None):
    d = dict(config=name, mode=mode, packets=pkts, seconds=secs, mpps=pkts / secs / 1e6, achieved_gbs=74 * pkts / secs / 1e9,
             frac_of_measured_hbm=74 * pkts / secs / 1e9 / PEAK)
    d.update(extra or {})
    print(json.dumps(d), flush=True)


def gen(dp, spec, n, first=0):
    hdr = torch.empty((n, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty(n, dtype=torch.int16, device="cuda")
    step = 1 << 22
    for o in range(0, n, step):
        m = min(step, n - o)
        dp.gen_packets(spec, first + o, m, hdr[o:].data_ptr(), ports[o:].data_ptr(), st)
    torch.cuda.synchronize()
    return hdr, ports


def cpu_rate(sc, hdr, ports, n):
    h, p = hdr[:n].cpu().numpy(), ports[:n].cpu().numpy().view(np.uint16)
    out = {}
    for thr in (1, os.cpu_count()):
        t = time.perf_counter(); sc.orc.classify(h, p, threads=thr); out[thr] = n / (time.perf_counter() - t) / 1e6
    return {"cpu_mpps_1thr": out[1], f"cpu_mpps_{os.cpu_count()}thr": out[os.cpu_count()], "cpu_sample_packets": n}


def cfg3():
    dp = sfc.Dataplane(0); sc = Scenario(dp)
    topo = sc.build_chain(8, 4)
    spec = config3_spec(n_flows=1 << 20, n_ports=topo["n_ports"])
    install_flows(sc, spec); dp.commit()
    B, NB = 4096, 4096
    n = B * NB
    hdr, ports = gen(dp, spec, n)
    res = torch.zeros(n, dtype=torch.int64, device="cuda")
    hp, pp, rp = hdr.data_ptr(), ports.data_ptr(), res.data_ptr()
    info = dict(table_bytes=dp.table_size, flows=1 << 20, batch=B, batches=NB)

    def per_batch(flags):
        for i in range(NB):
            dp.classify(hp + i * B * 64, pp + i * B * 2, rp + i * B * 8, B, flags, st)
    report("cfg3", "one launch per 4096-packet batch, serial", n, timed(lambda: per_batch(0)), info)
    report("cfg3", "one launch per 4096-packet batch, overlapped (PDL)", n, timed(lambda: per_batch(OVL)), info)
    try:
        g = torch.cuda.CUDAGraph()
        K = 512
        with torch.cuda.graph(g):
            s2 = torch.cuda.current_stream().cuda_stream
            for i in range(K):
                dp.classify(hp + i * B * 64, pp + i * B * 2, rp + i * B * 8, B, OVL, s2)
        report("cfg3", f"CUDA graph of {K} per-batch launches, replayed", B * K * 8, timed(lambda: [g.replay() for _ in range(8)]), info)
    except Exception as e:  # noqa: BLE001
        print(json.dumps(dict(config="cfg3", mode="CUDA graph", error=str(e)[:200])), flush=True)
    M = 1 << 20

    def big():
        for i in range(n // M):
            dp.classify(hp + i * M * 64, pp + i * M * 2, rp + i * M * 8, M, OVL, st)
    report("cfg3", "256 batches per launch (1 Mi packets), overlapped, 128 launches", 8 * n, timed(lambda: [big() for _ in range(8)]),
           dict(info, **cpu_rate(sc, hdr, ports, 1 << 20)))


def cfg4():
    dp = sfc.Dataplane(0); sc = Scenario(dp)
    spec = config4_spec()
    install_lpm_acl(sc, spec); dp.commit()
    n, M = 1 << 24, 1 << 20
    hdr, ports = gen(dp, spec, n)
    res = torch.zeros(n, dtype=torch.int64, device="cuda")
    hp, pp, rp = hdr.data_ptr(), ports.data_ptr(), res.data_ptr()

    def run():
        for i in range(n // M):
            dp.classify(hp + i * M * 64, pp + i * M * 2, rp + i * M * 8, M, OVL, st)
    info = dict(table_bytes=dp.table_size, prefixes=100000, acl_rules=1024)
    report("cfg4", "LPM v4/v6 + 1024-rule ACL, 128 x 1 Mi-packet launches", 8 * n, timed(lambda: [run() for _ in range(8)]),
           dict(info, **cpu_rate(sc, hdr, ports, 1 << 17)))
    # the same traffic with the ACL stage switched off (LPM cost alone)
    dp2 = sfc.Dataplane(0); sc2 = Scenario(dp2)
    for af, k in ((4, 80000), (6, 20000)):
        for p, ln, port in sfc.gen_prefixes(spec, af, k):
            sc2.lpm_add(af, p, ln, port)
    dp2.commit()

    def run2():
        for i in range(n // M):
            dp2.classify(hp + i * M * 64, pp + i * M * 2, rp + i * M * 8, M, OVL, st)
    report("cfg4", "LPM only (no ACL rules), 128 launches", 8 * n, timed(lambda: [run2() for _ in range(8)]), dict(table_bytes=dp2.table_size))


def cfg5():
    dp = sfc.Dataplane(0)
    spec = config3_spec(n_flows=1 << 24, n_ports=17)
    t = time.time()
    for first in range(0, 1 << 24, 1 << 20):
        k, a = sfc.gen_flows(spec, first, 1 << 20); dp.flow_add_bulk(k, a)
    dp.commit(); build_s = time.time() - t
    n, M = 1 << 24, 1 << 20
    hdr, ports = gen(dp, spec, n)
    res = torch.zeros(n, dtype=torch.int64, device="cuda")
    hp, pp, rp = hdr.data_ptr(), ports.data_ptr(), res.data_ptr()

    def run():
        for rep in range(4):
            for i in range(n // M):
                dp.classify(hp + i * M * 64, pp + i * M * 2, rp + i * M * 8, M, OVL | sfc.consts.F_COUNTERS, st)
    report("cfg5", "16 Mi-flow table, Zipf traffic, counters on, 64 x 1 Mi-packet launches, 1 GPU", 4 * n, timed(run, reps=2),
           dict(table_bytes=dp.table_size, build_seconds=build_s))


def churn():
    """Row f3: cost of flow churn on a 1 Mi-flow table — edits per second through the incremental commit,
    and what a classify launch queued right behind a commit pays."""
    dp = sfc.Dataplane(0)
    spec = config3_spec(n_flows=(1 << 20) + 400000, n_ports=17, v6_permille=0)
    keys, acts = sfc.gen_flows(spec, 0, (1 << 20) + 400000)
    dp.flow_add_bulk(keys[: 1 << 20], acts[: 1 << 20])
    t = time.perf_counter(); dp.commit(); full_s = time.perf_counter() - t
    M = 1 << 20
    hdr, ports = gen(dp, spec, M)
    res = torch.zeros(M, dtype=torch.int64, device="cuda")
    nxt, dead = 1 << 20, 0
    for edits in (64, 1024, 16384):
        times = []
        for rep in range(6):
            t0 = time.perf_counter()
            dp.flow_add_bulk(keys[nxt:nxt + edits // 2], acts[nxt:nxt + edits // 2]); nxt += edits // 2
            for k in keys[dead:dead + edits // 2]:
                dp.flow_del(k)
            dead += edits // 2
            t1 = time.perf_counter()
            dp.commit_stream(st)
            dp.classify(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), M, 0, st)
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            times.append((t1 - t0, t2 - t1))
        stage, commit = min(x[0] for x in times), min(x[1] for x in times)
        print(json.dumps(dict(config="churn", flows=1 << 20, edits_per_commit=edits, stage_seconds=stage, commit_plus_1Mi_classify_seconds=commit,
                              edits_per_second=edits / (stage + commit), full_rebuild_seconds=full_s,
                              incremental_commits=dp.stat("incremental_commits"), full_commits=dp.stat("full_commits"))), flush=True)


for name in sys.argv[1:] or ["cfg3", "cfg4", "cfg5"]:
    globals()[name]()
    torch.cuda.empty_cache()
