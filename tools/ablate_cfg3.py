"""Which ingredient of the config-3 traffic mix costs what: vary one knob at a time."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from common import Scenario, config3_spec, install_flows, sfc

M, POOL = 1 << 20, 8
st = torch.cuda.current_stream().cuda_stream
base = dict(n_flows=1 << 20, n_ports=17)
cases = {
    "cfg3 as specified": {},
    "no IPv6": dict(v6_permille=0),
    "no non-IP": dict(nonip_permille=0),
    "no VLAN tags": dict(vlan_permille=0),
    "uniform instead of Zipf": dict(zipf=0),
    "no misses": dict(hit_permille=1000),
    "v4 only, no tags, no non-IP": dict(v6_permille=0, vlan_permille=0, nonip_permille=0),
    "no chain rules": dict(_nochain=True),
    "3 hashes, load 80%": dict(_opts=dict(flow_hashes=3, flow_load_pct=80)),
    "3 hashes, load 60%": dict(_opts=dict(flow_hashes=3, flow_load_pct=60)),
    "2 hashes, load 45%": dict(_opts=dict(flow_load_pct=45)),
    "2 hashes, load 25%": dict(_opts=dict(flow_load_pct=25)),
}
if len(sys.argv) > 1:
    cases = {k: v for k, v in cases.items() if any(a in k for a in sys.argv[1:])}
for name, kw in cases.items():
    kw = dict(kw); nochain = kw.pop("_nochain", False); opts = kw.pop("_opts", {})
    dp = sfc.Dataplane(0); sc = Scenario(dp)
    for o, v in opts.items(): dp.set_option(o, v)
    if not nochain: sc.build_chain(8, 4)
    spec = config3_spec(**{**base, **kw})
    for f in range(0, 1 << 20, 1 << 18):
        k, a = sfc.gen_flows(spec, f, 1 << 18); dp.flow_add_bulk(k, a)
    dp.commit()
    hdr = torch.empty((POOL, M, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((POOL, M), dtype=torch.int16, device="cuda")
    res = torch.zeros((POOL, M), dtype=torch.int64, device="cuda")
    for b in range(POOL): dp.gen_packets(spec, b * M, M, hdr[b].data_ptr(), ports[b].data_ptr(), st)
    torch.cuda.synchronize()
    def run(k):
        for i in range(k):
            b = i % POOL
            dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), M, 2, st)
    run(200); torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(300); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 300)
    print(f"{name:32s} {best*1e3:7.2f} us/step {M/best/1e6:7.2f} Gpps  table {dp.table_size/1e6:.0f} MB", flush=True)
    del dp, sc, hdr, ports, res
