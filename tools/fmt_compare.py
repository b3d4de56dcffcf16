"""Flow table formats side by side on one B200: steady-state overlapped launches (us per Mi packets) and a
20-launch train, for table sizes from 1 K to 16 Mi flows; results of the two formats must be identical.
usage: python tools/fmt_compare.py [mode ...]   modes: cfg2 f2^18 f2^20 f2^20z cfg3 f2^22z big16mz"""
import importlib, json, os, statistics, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

sfc = importlib.import_module("dpu-operator_b200")
BATCH, POOL = 1 << 20, 8
OVL = sfc.consts.F_OVERLAP
st = torch.cuda.current_stream().cuda_stream


def spec_for(mode):
    if mode == "cfg2":
        return sfc.make_gen_spec(n_flows=1024, hit_permille=900), 1024
    if mode == "cfg3":
        nf = 1 << 20
        return sfc.make_gen_spec(n_flows=nf, n_ports=17, n_out_ports=17, hit_permille=947, nonip_permille=100,
                                 v6_permille=200, vlan_permille=250, zipf=1, n_macs=12), nf
    if mode.startswith("big16m"):
        nf = 1 << 24
        return sfc.make_gen_spec(n_flows=nf, n_ports=17, n_out_ports=17, hit_permille=947, nonip_permille=100,
                                 v6_permille=200, vlan_permille=250, zipf=int(mode.endswith("z")), n_macs=12), nf
    lg = int(mode[3:].rstrip("z")); nf = 1 << lg
    return sfc.make_gen_spec(n_flows=nf, hit_permille=950, zipf=int(mode.endswith("z"))), nf


def ev():
    return torch.cuda.Event(enable_timing=True)


for mode in sys.argv[1:] or ["cfg2", "f2^18", "f2^20", "cfg3"]:
    spec, nf = spec_for(mode)
    hdr = torch.empty((POOL, BATCH, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((POOL, BATCH), dtype=torch.int16, device="cuda")
    res = torch.zeros((POOL, BATCH), dtype=torch.int64, device="cuda")
    ref = None
    for fmt in (1, 2):
        dp = sfc.Dataplane(0)
        dp.set_option("flow_format", fmt)
        t0 = time.time()
        for f in range(0, nf, 1 << 20):
            k, a = sfc.gen_flows(spec, f, min(nf - f, 1 << 20)); dp.flow_add_bulk(k, a)
        dp.commit()
        build_s = time.time() - t0
        for b in range(POOL):
            dp.gen_packets(spec, b * BATCH, BATCH, hdr[b].data_ptr(), ports[b].data_ptr(), st)
        torch.cuda.synchronize()

        def run(k, flags=OVL):
            for i in range(k):
                b = i % POOL
                dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), BATCH, flags, st)

        for low in (0, 1):
            dp.set_option("low_regs", low)
            run(200); torch.cuda.synchronize()
            best = 1e9
            for rep in range(3):
                e0, e1 = ev(), ev()
                e0.record(); run(400); e1.record(); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) / 400 * 1e3)
            t20 = []
            for rep in range(9):
                torch.cuda.synchronize()
                e0, e1 = ev(), ev()
                e0.record(); run(20); e1.record(); torch.cuda.synchronize()
                t20.append(e0.elapsed_time(e1) / 20 * 1e3)
            run(8); torch.cuda.synchronize()
            cur = res.clone()
            if ref is None:
                ref = cur
            print(json.dumps(dict(mode=mode, flows=nf, fmt=fmt, low_regs=low, table_mb=round(dp.table_size / 1e6, 1), build_s=round(build_s, 1),
                                  us_per_Mi=round(best, 2), gpps=round(BATCH / best / 1e3, 2), frac=round(74 * BATCH / best / 1e3 / 6546.9, 3),
                                  train20_us=round(statistics.median(t20), 2), first_bucket=dp.stat("flow4_first_bucket"),
                                  same=bool(torch.equal(ref, cur)))), flush=True)
        del dp
    del hdr, ports, res, ref
    torch.cuda.empty_cache()
