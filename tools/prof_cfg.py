"""A few classify launches of one BASELINE config for ncu.  usage: python tools/prof_cfg.py cfg3|cfg5|cfg4 [launches]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from common import Scenario, config3_spec, config4_spec, install_flows, install_lpm_acl, sfc

which = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
launches = int(sys.argv[2]) if len(sys.argv) > 2 else 6
dp = sfc.Dataplane(0); sc = Scenario(dp)
if which == "cfg3":
    topo = sc.build_chain(8, 4)
    spec = config3_spec(n_flows=1 << 20, n_ports=topo["n_ports"])
    install_flows(sc, spec)
elif which == "cfg3u":  # 1 Mi IPv4 flows, IPv4-only traffic
    spec = sfc.make_gen_spec(n_flows=1 << 20, hit_permille=950)
    for f in range(0, 1 << 20, 1 << 18):
        k, a = sfc.gen_flows(spec, f, 1 << 18); dp.flow_add_bulk(k, a)
elif which.startswith("f2^"):  # 2^N IPv4 flows, uniform IPv4-only traffic
    nf = 1 << int(which[3:])
    spec = sfc.make_gen_spec(n_flows=nf, hit_permille=950)
    for f in range(0, nf, 1 << 18):
        k, a = sfc.gen_flows(spec, f, min(nf, 1 << 18)); dp.flow_add_bulk(k, a)
elif which == "cfg4":
    spec = config4_spec(); install_lpm_acl(sc, spec)
else:
    spec = config3_spec(n_flows=1 << 24, n_ports=17)
    for first in range(0, 1 << 24, 1 << 20):
        k, a = sfc.gen_flows(spec, first, 1 << 20); dp.flow_add_bulk(k, a)
dp.commit()
dp.set_option("flow_probe_mode", int(os.environ.get("SFC_PROBE", "-1")))
M = 1 << 20
st = torch.cuda.current_stream().cuda_stream
hdr = torch.empty((launches, M, 64), dtype=torch.uint8, device="cuda")
ports = torch.empty((launches, M), dtype=torch.int16, device="cuda")
res = torch.zeros((launches, M), dtype=torch.int64, device="cuda")
for i in range(launches):
    dp.gen_packets(spec, i * M, M, hdr[i].data_ptr(), ports[i].data_ptr(), st)
torch.cuda.synchronize()
flags = int(os.environ.get("SFC_FLAGS", "0"))
for rep in range(int(os.environ.get("SFC_REPS", "1"))):
    for i in range(launches):
        dp.classify(hdr[i].data_ptr(), ports[i].data_ptr(), res[i].data_ptr(), M, flags, st)
torch.cuda.synchronize()
print("done", which, dp.table_size)
