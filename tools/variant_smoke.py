"""A small pass through every kernel variant (plain / counters / ACL / multi-batch / static / patch / generator),
checked against the oracle at sizes that run in seconds (compute-sanitizer is not available on this pool).  usage: python tools/sanitize_smoke.py"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from common import Scenario, config3_spec, install_flows, install_lpm_acl, sfc

F = sfc.consts
st = torch.cuda.current_stream().cuda_stream
for with_acl in (False, True):
    dp = sfc.Dataplane(0); sc = Scenario(dp)
    topo = sc.build_chain(4, 2)
    spec = config3_spec(n_flows=3000, n_ports=topo["n_ports"] + 2, n_pfx4=400, n_pfx6=150, lpm_permille=600,
                        n_acl=60 if with_acl else 0, hit_permille=600)
    install_flows(sc, spec)
    if with_acl: install_lpm_acl(sc, spec)
    dp.commit()
    n = 3000 + 77
    hdr_h, port_h = sfc.gen_packets_host(spec, 0, n)
    want = sc.orc.classify(hdr_h, port_h)
    hdr = torch.empty((n, 64), dtype=torch.uint8, device="cuda"); ports = torch.empty(n, dtype=torch.int16, device="cuda")
    dp.gen_packets(spec, 0, n, hdr.data_ptr(), ports.data_ptr(), st)
    for flags in (0, F.F_COUNTERS, F.F_OVERLAP, F.F_OVERLAP | F.F_COUNTERS):
        res = torch.zeros(n, dtype=torch.int64, device="cuda")
        dp.classify(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), n, flags, st)
        torch.cuda.synchronize()
        assert np.array_equal(res.cpu().numpy().view(np.uint64), want), ("classify", with_acl, flags)
        res.zero_()
        cuts = [0, 1, 300, 556, 812, 1500, n]  # ragged batches, some not a multiple of the tile
        batches = [(hdr.data_ptr() + a * 64, ports.data_ptr() + a * 2, res.data_ptr() + a * 8, b - a) for a, b in zip(cuts, cuts[1:])]
        dp.classify_batches(batches, flags, st)
        torch.cuda.synchronize()
        assert np.array_equal(res.cpu().numpy().view(np.uint64), want), ("batches", with_acl, flags)
    assert np.array_equal(dp.classify_host(hdr_h, port_h), want)
    # incremental commit: patch kernel
    k, a = sfc.gen_flows(spec, 0, 3000)
    for i in range(0, 200, 2): dp.flow_del(k[i]); sc.orc.flow_del(k[i])
    dp.commit_stream(st)
    res = torch.zeros(n, dtype=torch.int64, device="cuda")
    dp.classify(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), n, 0, st); torch.cuda.synchronize()
    assert np.array_equal(res.cpu().numpy().view(np.uint64), sc.orc.classify(hdr_h, port_h)), ("after patch", with_acl)
    # static entry point
    img = dp.table_export()
    if img.size <= sfc.STATIC_AREA_BYTES - 256:
        res.zero_()
        sfc.classify_static(0, img, hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), n, st); torch.cuda.synchronize()
        assert np.array_equal(res.cpu().numpy().view(np.uint64), sc.orc.classify(hdr_h, port_h)), ("static", with_acl)
    print("variant set ok, acl =", with_acl, flush=True)
print("variant smoke ok")
