"""Repeat every compiled kernel variant many times and compare full result arrays (race hunt)."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from common import Oracle, sfc

BATCH, POOL = 1 << 20, 8
VARIANTS = [(0, 0), (128, 2), (128, 3), (256, 2), (256, 3), (256, 4), (512, 2), (512, 3), (768, 2)]
dp = sfc.Dataplane(0)
spec = sfc.make_gen_spec(n_flows=1024, hit_permille=900)
k, a = sfc.gen_flows(spec, 0, 1024); dp.flow_add_bulk(k, a); dp.commit()
orc = Oracle(); orc.flow_add(k, a, 0)
st = torch.cuda.current_stream().cuda_stream
hdr = torch.empty((POOL, BATCH, 64), dtype=torch.uint8, device="cuda")
ports = torch.empty((POOL, BATCH), dtype=torch.int16, device="cuda")
res = torch.zeros((POOL, BATCH), dtype=torch.int64, device="cuda")
for b in range(POOL):
    dp.gen_packets(spec, b * BATCH, BATCH, hdr[b].data_ptr(), ports[b].data_ptr(), st)
torch.cuda.synchronize()
want = torch.stack([torch.from_numpy(orc.classify(hdr[b].cpu().numpy(), ports[b].cpu().numpy().view(np.uint16), threads=16).view(np.int64)) for b in range(POOL)]).cuda()
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 40
for tile, stages in VARIANTS:
    dp.set_option("tune_tile", tile); dp.set_option("tune_stages", stages)
    bad_total = 0
    for it in range(iters):
        res.zero_()
        for i in range(POOL * 4):
            b = i % POOL
            dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), BATCH, int(os.environ.get("SFC_FLAGS", "0")), st)
        torch.cuda.synchronize()
        bad = (res != want).nonzero()
        if bad.numel():
            bad_total += bad.shape[0]
            if bad_total == bad.shape[0]:
                rows = bad[:12].cpu().tolist()
                print(f"  variant {tile},{stages} iter {it}: {bad.shape[0]} mismatches, first {rows}")
                for b_, i_ in rows[:4]:
                    print(f"    got {res[b_, i_].item() & (2**64-1):#018x} want {want[b_, i_].item() & (2**64-1):#018x}")
    print(f"variant tile={tile} stages={stages}: {bad_total} mismatching results over {iters} x {POOL*4} launches", flush=True)
