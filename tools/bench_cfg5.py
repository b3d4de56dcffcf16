"""BASELINE.json configs[4]: a 16 Mi-flow table built on rank 0, broadcast over NCCL, 2^28 packets sharded
contiguously across the ranks, per-action/per-stage counters all-reduced at the end.  Launch with
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_cfg5.py
(or plain `python tools/bench_cfg5.py` for one GPU).  Prints one JSON line on rank 0."""
import importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import torch.distributed as dist
from common import Oracle, config3_spec, sfc

group = importlib.import_module("dpu-operator_b200.group")
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
N_FLOWS, TOTAL, M = 1 << 24, int(os.environ.get("CFG5_PACKETS", 1 << 28)), 1 << 20
dp = sfc.Dataplane(local)
spec = config3_spec(n_flows=N_FLOWS, n_ports=17)
build_s = 0.0
if rank == 0:
    t = time.perf_counter()
    for first in range(0, N_FLOWS, 1 << 20):
        k, a = sfc.gen_flows(spec, first, 1 << 20); dp.flow_add_bulk(k, a)
    dp.commit(); build_s = time.perf_counter() - t
bcast_ms = 0.0
if world > 1:
    group.broadcast_table(dp, rank, world)        # first call pays NCCL initialisation
    bcast_ms = group.broadcast_table(dp, rank, world)  # the steady-state cost of replicating an image
lo, hi = group.shard_range(TOTAL, rank, world)
n = hi - lo
st = torch.cuda.current_stream().cuda_stream
hdr = torch.empty((n, 64), dtype=torch.uint8, device="cuda")
ports = torch.empty(n, dtype=torch.int16, device="cuda")
res = torch.zeros(n, dtype=torch.int64, device="cuda")
for o in range(0, n, 1 << 22):
    dp.gen_packets(spec, lo + o, min(1 << 22, n - o), hdr[o:].data_ptr(), ports[o:].data_ptr(), st)
torch.cuda.synchronize()
flags = sfc.consts.F_OVERLAP | sfc.consts.F_COUNTERS


def run():
    for o in range(0, n, M):
        m = min(M, n - o)
        dp.classify(hdr[o:].data_ptr(), ports[o:].data_ptr(), res[o:].data_ptr(), m, flags, st)


run(); torch.cuda.synchronize()  # warm-up pass
dp.counters_reset()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); run(); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
if world > 1:
    t = torch.tensor([ms], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
reduce_ms = group.reduce_counters(dp, world) if world > 1 else 0.0
cnt = dp.counters_read(16)
# property check on this rank's shard: a flow hit names a flow whose key is the packet's key
idx = torch.randint(0, n, (2000,), device="cuda")
got = res[idx].cpu().numpy().view(np.uint64); hs = hdr[idx].cpu().numpy()
act, nf, port, stage, fid = sfc.result_fields(got)
ok = 0
for i in range(len(got)):
    if stage[i] != sfc.consts.STAGE_FLOW:
        continue
    k, a = sfc.gen_flows(spec, int(fid[i]), 1); p = Oracle.parse(hs[i]); w = 4 if k[0]["ip_ver"] == 4 else 16
    assert p["src"][:w] == bytes(k[0]["src"][:w]) and p["dst"][:w] == bytes(k[0]["dst"][:w]) and p["sport"] == k[0]["sport"] and act[i] == a[0]["action"]
    ok += 1
if rank == 0:
    assert int(cnt[0] + cnt[1] + cnt[2]) == TOTAL, (cnt[:3], TOTAL)
    print(json.dumps(dict(config="cfg5", n_gpus=world, flows=N_FLOWS, packets=TOTAL, table_bytes=dp.table_size, seconds=ms * 1e-3,
                          mpps=TOTAL / (ms * 1e-3) / 1e6, per_gpu_gbs=74 * (TOTAL / world) / (ms * 1e-3) / 1e9,
                          build_seconds=build_s, broadcast_ms=bcast_ms, counter_allreduce_ms=reduce_ms,
                          counters=dict(to_cp=int(cnt[0]), forward=int(cnt[1]), drop=int(cnt[2]), flow_hits=int(cnt[6])),
                          checked_flow_hits=ok)), flush=True)
if world > 1:
    dist.destroy_process_group()
