"""configs[3] (LPM + ACL) under different grid policies.  usage: python tools/tune_cfg4.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from common import Scenario, config4_spec, install_lpm_acl, sfc

dp = sfc.Dataplane(0); sc = Scenario(dp)
spec = config4_spec(); install_lpm_acl(sc, spec); dp.commit()
M, POOL = 1 << 20, 8
st = torch.cuda.current_stream().cuda_stream
hdr = torch.empty((POOL, M, 64), dtype=torch.uint8, device="cuda")
ports = torch.empty((POOL, M), dtype=torch.int16, device="cuda")
res = torch.zeros((POOL, M), dtype=torch.int64, device="cuda")
for b in range(POOL):
    dp.gen_packets(spec, b * M, M, hdr[b].data_ptr(), ports[b].data_ptr(), st)
torch.cuda.synchronize()
ref = None
dp.set_option("tune_carveout", int(os.environ.get("SFC_CARVEOUT", "-1")))
for flags, ctas in ((0, 0), (2, 0)):
    dp.set_option("tune_ctas", ctas)
    def run(k):
        for i in range(k):
            b = i % POOL
            dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), M, flags, st)
    run(20); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(100); e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 10
    cur = res[0].clone(); ref = cur if ref is None else ref
    print(f"flags={flags} ctas/SM={ctas or ('max' if not flags else 1)}: {us:7.2f} us per Mi  {M/us/1e3:6.2f} Gpps same={bool(torch.equal(ref, cur))}", flush=True)
