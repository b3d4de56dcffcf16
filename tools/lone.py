"""How long ONE launch takes, a train of 20, and a long train — per (stages, CTAs per SM, overlap) variant.
Needs the experiment build (make -C dpu-operator_b200/csrc tune) and B200SFC_LIB pointing at it.
usage: B200SFC_LIB=dpu-operator_b200/libb200sfc_tune.so python tools/lone.py [cfg2|empty|f2^18 ...]"""
import importlib, json, os, statistics, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

sfc = importlib.import_module("dpu-operator_b200")
BATCH, POOL = 1 << 20, 8
OVL = sfc.consts.F_OVERLAP
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def setup(mode):
    dp = sfc.Dataplane(0)
    if mode == "empty":
        spec = sfc.make_gen_spec(n_flows=1024, hit_permille=900)
    elif mode == "cfg2":
        spec = sfc.make_gen_spec(n_flows=1024, hit_permille=900)
        k, a = sfc.gen_flows(spec, 0, 1024); dp.flow_add_bulk(k, a)
    elif mode.startswith("f2^"):
        lg = int(mode[3:].rstrip("z")); nf = 1 << lg
        spec = sfc.make_gen_spec(n_flows=nf, hit_permille=950, zipf=int(mode.endswith("z")))
        for f in range(0, nf, 1 << 18):
            k, a = sfc.gen_flows(spec, f, min(nf, 1 << 18)); dp.flow_add_bulk(k, a)
    dp.commit()
    return dp, spec


def ev():
    return torch.cuda.Event(enable_timing=True)


for mode in sys.argv[1:] or ["cfg2"]:
    dp, spec = setup(mode)
    hdr = torch.empty((POOL, BATCH, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((POOL, BATCH), dtype=torch.int16, device="cuda")
    res = torch.zeros((POOL, BATCH), dtype=torch.int64, device="cuda")
    for b in range(POOL):
        dp.gen_packets(spec, b * BATCH, BATCH, hdr[b].data_ptr(), ports[b].data_ptr(), st)
    torch.cuda.synchronize()

    def run(k, flags):
        for i in range(k):
            b = i % POOL
            dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), BATCH, flags, st)

    variants = os.environ.get("LONE_VARIANTS", "0x0,256x2,256x3,256x4,256x6,256x8,128x4").split(",")
    ctas_list = [int(x) for x in os.environ.get("LONE_CTAS", "0,1,2,3").split(",")]
    for v in variants:
        tile, stages = map(int, v.split("x"))
        for ctas in ctas_list:
            for flags in (OVL, 0):
                dp.set_option("tune_tile", tile); dp.set_option("tune_stages", stages); dp.set_option("tune_ctas", ctas)
                try:
                    run(50, flags); torch.cuda.synchronize()
                except Exception as e:  # noqa: BLE001
                    print(json.dumps(dict(mode=mode, variant=v, ctas=ctas, error=str(e)[:120])), flush=True)
                    continue
                lone, t20 = [], []
                for rep in range(9):
                    flush.fill_(rep); torch.cuda.synchronize()
                    e0, e1 = ev(), ev()
                    e0.record(); run(1, flags); e1.record(); torch.cuda.synchronize()
                    lone.append(e0.elapsed_time(e1) * 1e3)
                for rep in range(9):
                    torch.cuda.synchronize()
                    e0, e1 = ev(), ev()
                    e0.record(); run(20, flags); e1.record(); torch.cuda.synchronize()
                    t20.append(e0.elapsed_time(e1) * 1e3 / 20)
                run(200, flags); torch.cuda.synchronize()
                e0, e1 = ev(), ev()
                e0.record(); run(1000, flags); e1.record(); torch.cuda.synchronize()
                long_us = e0.elapsed_time(e1) * 1e3 / 1000
                print(json.dumps(dict(mode=mode, variant=v, ctas=ctas, overlap=bool(flags), lone_us=round(statistics.median(lone), 2),
                                      train20_us=round(statistics.median(t20), 2), long_us=round(long_us, 2),
                                      frac20=round(74 * BATCH / statistics.median(t20) / 1e3 / 6546.9, 3))), flush=True)
    del hdr, ports, res, dp
