"""One classify launch over 16 Mi packets (configs[1] table), not overlapped with anything: the case in which ramp-up
and tail are amortised, for an ncu capture of the streaming rate.  usage: python tools/prof_big.py [launches]"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

sfc = importlib.import_module("dpu-operator_b200")
launches = int(sys.argv[1]) if len(sys.argv) > 1 else 3
N = 1 << 24
dp = sfc.Dataplane(0)
spec = sfc.make_gen_spec(n_flows=1024, hit_permille=900)
dp.flow_add_bulk(*sfc.gen_flows(spec, 0, 1024))
dp.commit()
st = torch.cuda.current_stream().cuda_stream
hdr = torch.empty((N, 64), dtype=torch.uint8, device="cuda")
ports = torch.empty(N, dtype=torch.int16, device="cuda")
res = torch.zeros(N, dtype=torch.int64, device="cuda")
for o in range(0, N, 1 << 22):
    dp.gen_packets(spec, o, 1 << 22, hdr.data_ptr() + o * 64, ports.data_ptr() + o * 2, st)
torch.cuda.synchronize()
e = [torch.cuda.Event(enable_timing=True) for _ in range(launches + 1)]
e[0].record()
for i in range(launches):
    dp.classify(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), N, 0, st)
    e[i + 1].record()
torch.cuda.synchronize()
for i in range(launches):
    ms = e[i].elapsed_time(e[i + 1])
    print(f"launch {i}: {ms * 1e3:.1f} us for 16 Mi packets = {N / ms / 1e6:.1f} Gpps = {74 * N / ms / 1e6:.0f} GB/s algorithmic")
