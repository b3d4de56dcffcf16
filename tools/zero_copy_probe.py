"""Experiment: let the classify kernel read packet headers straight from pinned host memory (UVA) and write the
results straight back, instead of staging through device buffers.  usage: python tools/zero_copy_probe.py"""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
sfc = importlib.import_module("dpu-operator_b200")
dp = sfc.Dataplane(0)
spec = sfc.make_gen_spec(n_flows=1024, hit_permille=900)
k, a = sfc.gen_flows(spec, 0, 1024); dp.flow_add_bulk(k, a); dp.commit()
N = 1 << 20
hdr_np, port_np = sfc.gen_packets_host(spec, 0, N)
hdr = torch.from_numpy(np.array(hdr_np)).pin_memory(); ports = torch.from_numpy(np.array(port_np).view(np.int16)).pin_memory()
res = torch.zeros(N, dtype=torch.int64).pin_memory()
ref = dp.classify_host(hdr_np, port_np)
st = torch.cuda.current_stream().cuda_stream
for mode in ("results to device", "results to pinned host"):
    out = torch.zeros(N, dtype=torch.int64, device="cuda") if mode == "results to device" else res
    for flags in (0, 2):
        def run(n):
            for _ in range(n):
                dp.classify(hdr.data_ptr(), ports.data_ptr(), out.data_ptr(), N, flags, st)
        run(2); torch.cuda.synchronize()
        t = time.perf_counter(); run(10); torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 10
        ok = np.array_equal((out.cpu() if out.is_cuda else out).numpy().view(np.uint64), ref)
        print(f"zero-copy, {mode}, flags={flags}: {dt*1e3:.3f} ms per Mi packets = {N/dt/1e6:.0f} Mpps, {66*N/dt/1e9:.1f} GB/s in, same={ok}", flush=True)
t = time.perf_counter()
for _ in range(10): dp.classify_host_ptr(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), N, 0)
dt = (time.perf_counter() - t) / 10
print(f"sfc_classify_host (staged): {dt*1e3:.3f} ms = {N/dt/1e6:.0f} Mpps")
