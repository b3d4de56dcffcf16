import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from common import Scenario, config4_spec, sfc
spec = config4_spec()
for lr in (0, 1):
    dp = sfc.Dataplane(0); sc = Scenario(dp)
    dp.set_option("low_regs", lr)
    for af, k in ((4, 80000), (6, 20000)):
        for p, ln, port in sfc.gen_prefixes(spec, af, k): sc.lpm_add(af, p, ln, port)
    dp.commit()
    M, POOL = 1 << 20, 8
    st = torch.cuda.current_stream().cuda_stream
    hdr = torch.empty((POOL, M, 64), dtype=torch.uint8, device="cuda"); ports = torch.empty((POOL, M), dtype=torch.int16, device="cuda")
    res = torch.zeros((POOL, M), dtype=torch.int64, device="cuda")
    for b in range(POOL): dp.gen_packets(spec, b * M, M, hdr[b].data_ptr(), ports[b].data_ptr(), st)
    def run(k):
        for i in range(k):
            b = i % POOL; dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), M, 2, st)
    run(50); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(200); e1.record(); torch.cuda.synchronize()
    print("low_regs", lr, round(e0.elapsed_time(e1) * 5, 2), "us per Mi")
