"""Probe policy vs traffic skew: lazy / lazy+L1 / eager for the sector tables (and the indexed table) on uniform and
Zipf traffic.  usage: python tools/probe_modes.py [mode ...]   (modes as in tools/fmt_compare.py)"""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch

sfc = importlib.import_module("dpu-operator_b200")
BATCH, POOL = 1 << 20, 8
OVL = sfc.consts.F_OVERLAP
st = torch.cuda.current_stream().cuda_stream


def spec_for(mode):
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


for mode in sys.argv[1:] or ["f2^20", "f2^20z", "cfg3"]:
    spec, nf = spec_for(mode)
    hdr = torch.empty((POOL, BATCH, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((POOL, BATCH), dtype=torch.int16, device="cuda")
    res = torch.zeros((POOL, BATCH), dtype=torch.int64, device="cuda")
    ref = None
    for fmt in (1, 2):
        if fmt == 2 and nf > (1 << 22):
            continue
        dp = sfc.Dataplane(0)
        dp.set_option("flow_format", fmt)
        for f in range(0, nf, 1 << 20):
            k, a = sfc.gen_flows(spec, f, min(nf - f, 1 << 20)); dp.flow_add_bulk(k, a)
        dp.commit()
        for b in range(POOL):
            dp.gen_packets(spec, b * BATCH, BATCH, hdr[b].data_ptr(), ports[b].data_ptr(), st)
        torch.cuda.synchronize()

        def run(k):
            for i in range(k):
                b = i % POOL
                dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), BATCH, OVL, st)

        for pm in ((0, 2, 1, 3) if fmt == 1 else (0, 2)):
            for low in (0, 1):
                dp.set_option("flow_probe_mode", pm); dp.set_option("low_regs", low)
                run(200); torch.cuda.synchronize()
                best = 1e9
                for rep in range(3):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(); run(400); e1.record(); torch.cuda.synchronize()
                    best = min(best, e0.elapsed_time(e1) / 400 * 1e3)
                run(8); torch.cuda.synchronize()
                cur = res.clone()
                if ref is None:
                    ref = cur
                print(json.dumps(dict(mode=mode, flows=nf, fmt=fmt, probe_mode=pm, low_regs=low, table_mb=round(dp.table_size / 1e6, 1),
                                      us_per_Mi=round(best, 2), frac=round(74 * BATCH / best / 1e3 / 6546.9, 3), same=bool(torch.equal(ref, cur)))), flush=True)
        del dp
    del hdr, ports, res, ref
    torch.cuda.empty_cache()
