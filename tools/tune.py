"""Sweep the compiled (tile, stages) variants of the classify kernel on a B200 (experiments only).
usage: python tools/tune.py [mode ...]   modes: empty cfg2 cfg3 big16m"""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

sfc = importlib.import_module("dpu-operator_b200")
BATCH, POOL = 1 << 20, 8
VARIANTS = [(0, 0), (64, 4), (128, 2), (128, 3), (256, 2), (256, 3), (256, 4), (512, 2), (512, 3), (768, 2)]


def setup(mode):
    dp = sfc.Dataplane(0)
    dp.set_option("flow_slots", int(os.environ.get("SFC_SLOTS", "0")))
    dp.set_option("flow_load_pct", int(os.environ.get("SFC_LOAD", "40")))
    dp.set_option("low_regs", int(os.environ.get("SFC_LOWREGS", "-1")))
    if mode == "empty":
        spec = sfc.make_gen_spec(n_flows=1024, hit_permille=900)
    elif mode == "cfg2":
        spec = sfc.make_gen_spec(n_flows=1024, hit_permille=900)
        k, a = sfc.gen_flows(spec, 0, 1024); dp.flow_add_bulk(k, a)
    elif mode == "cfg3":
        spec = sfc.make_gen_spec(n_flows=1 << 20, n_ports=17, n_out_ports=17, hit_permille=947, nonip_permille=100,
                                 v6_permille=200, vlan_permille=250, zipf=1, n_macs=12)
        for f in range(0, 1 << 20, 1 << 18):
            k, a = sfc.gen_flows(spec, f, 1 << 18); dp.flow_add_bulk(k, a)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
    elif mode == "cfg3u":  # 1M flows, IPv4 only, uniform
        spec = sfc.make_gen_spec(n_flows=1 << 20, hit_permille=950)
        for f in range(0, 1 << 20, 1 << 18):
            k, a = sfc.gen_flows(spec, f, 1 << 18); dp.flow_add_bulk(k, a)
    elif mode.startswith("f2^"):  # 2^N IPv4 flows, uniform or Zipf ("f2^18z") traffic: the L2 capacity curve
        lg = int(mode[3:].rstrip("z")); nf = 1 << lg
        spec = sfc.make_gen_spec(n_flows=nf, hit_permille=950, zipf=int(mode.endswith("z")))
        for f in range(0, nf, 1 << 18):
            k, a = sfc.gen_flows(spec, f, min(nf, 1 << 18)); dp.flow_add_bulk(k, a)
    elif mode == "big16m":
        spec = sfc.make_gen_spec(n_flows=1 << 24, hit_permille=950)
        for f in range(0, 1 << 24, 1 << 20):
            k, a = sfc.gen_flows(spec, f, 1 << 20); dp.flow_add_bulk(k, a)
    t = time.time(); dp.commit(); print(f"# {mode}: commit {time.time()-t:.1f}s table {dp.table_size/1e6:.1f} MB", flush=True)
    return dp, spec


FLAGS = int(os.environ.get("SFC_FLAGS", "0"))
for mode in sys.argv[1:] or ["empty", "cfg2"]:
    dp, spec = setup(mode)
    st = torch.cuda.current_stream().cuda_stream
    hdr = torch.empty((POOL, BATCH, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((POOL, BATCH), dtype=torch.int16, device="cuda")
    res = torch.zeros((POOL, BATCH), dtype=torch.int64, device="cuda")
    for b in range(POOL):
        dp.gen_packets(spec, b * BATCH, BATCH, hdr[b].data_ptr(), ports[b].data_ptr(), st)
    torch.cuda.synchronize()
    ref = None
    pm = int(os.environ.get("SFC_PROBE", "-1"))
    dp.set_option("flow_probe_mode", pm)
    for tile, stages in [tuple(map(int, v.split("x"))) for v in os.environ.get("SFC_VARIANTS", "0x0,128x2,128x3,256x2,256x3,512x2").split(",")]:
        dp.set_option("tune_tile", tile); dp.set_option("tune_stages", stages)
        dp.set_option("tune_carveout", int(os.environ.get("SFC_CARVEOUT", "-1")))
        dp.set_option("tune_ctas", int(os.environ.get("SFC_CTAS", "0")))
        def run(k):
            for i in range(k):
                b = i % POOL
                dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), BATCH, FLAGS, st)
        run(300); torch.cuda.synchronize()
        best = 1e9
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); run(400); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 400)
        cur = res[0].clone()
        if ref is None: ref = cur
        same = bool(torch.equal(ref, cur))
        print(f"{mode:7s} tile={tile:4d} stages={stages} : {best*1e3:7.2f} us/step  {BATCH/best/1e6:7.2f} Gpps  frac={74*BATCH/best/1e6/6546.9:.3f} same={same}", flush=True)
    del hdr, ports, res, dp
