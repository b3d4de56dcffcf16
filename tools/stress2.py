"""Race hunt over the variants tools/stress.py does not reach: the one-stage ACL kernel, the 40-register kernels,
the multi-batch kernel and the counters variants, all under overlapped launches.  Every launch's full result array
is compared with the oracle's.  usage: python tools/stress2.py [iterations]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from common import Scenario, config3_spec, install_flows, install_lpm_acl, sfc

F = sfc.consts
N, POOL = 1 << 18, 6
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20
st = torch.cuda.current_stream().cuda_stream
for name, with_acl, low_regs in (("acl one-stage", True, -1), ("plain 51 regs", False, 0), ("plain 40 regs", False, 1)):
    dp = sfc.Dataplane(0); sc = Scenario(dp)
    dp.set_option("low_regs", low_regs)
    topo = sc.build_chain(8, 3)
    spec = config3_spec(n_flows=200000, n_ports=topo["n_ports"] + 2, n_pfx4=5000, n_pfx6=2000, lpm_permille=600,
                        n_acl=200 if with_acl else 0, hit_permille=700)
    install_flows(sc, spec)
    install_lpm_acl(sc, spec) if with_acl else None
    if not with_acl:
        for af, k in ((4, 5000), (6, 2000)):
            for p, ln, port in sfc.gen_prefixes(spec, af, k): sc.lpm_add(af, p, ln, port)
    dp.commit()
    hdr = torch.empty((POOL, N, 64), dtype=torch.uint8, device="cuda"); ports = torch.empty((POOL, N), dtype=torch.int16, device="cuda")
    res = torch.zeros((POOL, N), dtype=torch.int64, device="cuda")
    for b in range(POOL): dp.gen_packets(spec, b * N, N, hdr[b].data_ptr(), ports[b].data_ptr(), st)
    torch.cuda.synchronize()
    want = torch.stack([torch.from_numpy(sc.orc.classify(hdr[b].cpu().numpy(), ports[b].cpu().numpy().view(np.uint16), threads=16).view(np.int64))
                        for b in range(POOL)]).cuda()
    cuts = [0, 4096, 4096 + 255, 70000, 70000 + 513, 200000, N]
    for mode in ("classify", "classify+counters", "batches", "batches+counters"):
        flags = F.F_OVERLAP | (F.F_COUNTERS if "counters" in mode else 0)
        bad_total = 0
        for it in range(iters):
            res.zero_()
            for i in range(POOL * 6):
                b = i % POOL
                if mode.startswith("batches"):
                    dp.classify_batches([(hdr[b].data_ptr() + a * 64, ports[b].data_ptr() + a * 2, res[b].data_ptr() + a * 8, e - a)
                                         for a, e in zip(cuts, cuts[1:])], flags, st)
                else:
                    dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), N, flags, st)
            torch.cuda.synchronize()
            bad_total += int((res != want).sum())
        print(f"{name:14s} {mode:18s}: {bad_total} mismatching results over {iters} x {POOL * 6} launches", flush=True)
