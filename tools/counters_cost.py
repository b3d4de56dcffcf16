import importlib, os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from common import config3_spec, sfc
dp = sfc.Dataplane(0)
nf = 1 << 22
spec = config3_spec(n_flows=nf, n_ports=17)
for first in range(0, nf, 1 << 20):
    k, a = sfc.gen_flows(spec, first, 1 << 20); dp.flow_add_bulk(k, a)
dp.commit()
M, POOL = 1 << 20, 8
st = torch.cuda.current_stream().cuda_stream
hdr = torch.empty((POOL, M, 64), dtype=torch.uint8, device="cuda"); ports = torch.empty((POOL, M), dtype=torch.int16, device="cuda")
res = torch.zeros((POOL, M), dtype=torch.int64, device="cuda")
for b in range(POOL): dp.gen_packets(spec, b * M, M, hdr[b].data_ptr(), ports[b].data_ptr(), st)
torch.cuda.synchronize()
for flags in (2, 3):
    def run(k):
        for i in range(k):
            b = i % POOL
            dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), M, flags, st)
    run(50); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(200); e1.record(); torch.cuda.synchronize()
    print("flags", flags, round(e0.elapsed_time(e1) * 5, 2), "us per Mi")
