"""configs[1] (1024 flows, table meant to live in L1): does a denser table pay?  Variants of load factor, bucket choices
and slots per bucket, steady-state overlapped launches.  usage: python tools/small_table.py"""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

sfc = importlib.import_module("dpu-operator_b200")
BATCH, POOL = 1 << 20, 8
OVL = sfc.consts.F_OVERLAP
st = torch.cuda.current_stream().cuda_stream
spec = sfc.make_gen_spec(n_flows=1024, hit_permille=900)
hdr = torch.empty((POOL, BATCH, 64), dtype=torch.uint8, device="cuda")
ports = torch.empty((POOL, BATCH), dtype=torch.int16, device="cuda")
res = torch.zeros((POOL, BATCH), dtype=torch.int64, device="cuda")
ref = None
for name, opts in (("default: 2 choices, 40 % load", {}), ("2 choices, 45 %", dict(flow_load_pct=45)), ("3 choices, 60 %", dict(flow_hashes=3, flow_load_pct=60)),
                   ("3 choices, 80 %", dict(flow_hashes=3, flow_load_pct=80)), ("two-slot buckets, 80 %", dict(flow_slots=2)),
                   ("2 choices, 25 %", dict(flow_load_pct=25)), ("indexed format", dict(flow_format=2))):
    dp = sfc.Dataplane(0)
    for k, v in opts.items():
        dp.set_option(k, v)
    dp.flow_add_bulk(*sfc.gen_flows(spec, 0, 1024))
    dp.commit()
    for b in range(POOL):
        dp.gen_packets(spec, b * BATCH, BATCH, hdr[b].data_ptr(), ports[b].data_ptr(), st)
    torch.cuda.synchronize()

    def run(k):
        for i in range(k):
            b = i % POOL
            dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), BATCH, OVL, st)

    for pm in (-1, 1, 3, 0, 2):
        dp.set_option("flow_probe_mode", pm)
        run(200); torch.cuda.synchronize()
        best = 1e9
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); run(400); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 400 * 1e3)
        run(8); torch.cuda.synchronize()
        cur = res.clone()
        ref = cur if ref is None else ref
        print(json.dumps(dict(table=name, probe_mode=pm, table_bytes=dp.table_size, us_per_Mi=round(best, 2),
                              frac=round(74 * BATCH / best / 1e3 / 6546.9, 3), first_bucket=dp.stat("flow4_first_bucket"), same=bool(torch.equal(ref, cur)))), flush=True)
    del dp
