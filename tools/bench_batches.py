"""configs[2] as worded — 1 Mi flows, 4096-packet batches — three ways of submitting the same work:
one sfc_classify per batch, the same captured in a CUDA graph, and sfc_classify_batches (256 batches per launch).
usage: python tools/bench_batches.py [batch_size]"""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from common import Scenario, config3_spec, install_flows, sfc

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
NB, POOL = (1 << 20) // B, 8
dp = sfc.Dataplane(0); sc = Scenario(dp)
topo = sc.build_chain(8, 4)
spec = config3_spec(n_flows=1 << 20, n_ports=topo["n_ports"])
install_flows(sc, spec); dp.commit()
s = torch.cuda.Stream(); st = s.cuda_stream
N = NB * B
hdr = torch.empty((POOL, N, 64), dtype=torch.uint8, device="cuda")
ports = torch.empty((POOL, N), dtype=torch.int16, device="cuda")
res = torch.zeros((POOL, N), dtype=torch.int64, device="cuda")
for p in range(POOL):
    dp.gen_packets(spec, p * N, N, hdr[p].data_ptr(), ports[p].data_ptr(), 0)
torch.cuda.synchronize()
lists = []
for p in range(POOL):
    arr = np.zeros(NB, dtype=sfc.BATCH_DTYPE)
    for i in range(NB):
        arr[i] = (hdr[p].data_ptr() + i * B * 64, ports[p].data_ptr() + i * B * 2, res[p].data_ptr() + i * B * 8, B, 0)
    lists.append(arr)


def timed(fn, reps):
    fn(2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        e0.record(s); fn(reps); e1.record(s)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps  # ms per 1 Mi packets


def per_batch(reps):
    for r in range(reps):
        a = lists[r % POOL]
        for i in range(NB):
            dp.classify(int(a["hdr"][i]), int(a["in_port"][i]), int(a["result"][i]), B, sfc.consts.F_OVERLAP, st)


def one_call(reps):
    for r in range(reps):
        dp.classify_batches(lists[r % POOL], sfc.consts.F_OVERLAP, st)


def big(reps):
    for r in range(reps):
        p = r % POOL
        dp.classify(hdr[p].data_ptr(), ports[p].data_ptr(), res[p].data_ptr(), N, sfc.consts.F_OVERLAP, st)


out = {"workload": f"configs[2]: 1Mi flows, {B}-packet batches, 4-NF chain", "batch": B, "batches_per_step": NB}
for name, fn, reps in (("one_launch_per_batch", per_batch, 8), ("classify_batches", one_call, 400), ("one_contiguous_launch", big, 400)):
    ms = timed(fn, reps)
    out[name] = {"us_per_Mi_packets": round(ms * 1e3, 2), "Gpps": round(N / ms / 1e6, 2), "roofline_frac": round(74 * N / ms / 1e6 / 6546.9, 3)}
ref = res[0].clone(); res[0].zero_(); big(1); torch.cuda.synchronize()
out["same_results"] = bool(torch.equal(ref, res[0]))
print(json.dumps(out))
