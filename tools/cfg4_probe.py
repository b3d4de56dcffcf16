"""Why is configs[4] slow in bench.py?  One 16 Mi-flow table, then the timed pass with: counters on/off, chain rules
in the table or not, 2^24 / 2^26 / 2^28 packets.  usage: python tools/cfg4_probe.py"""
import importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch
import bench

sfc = importlib.import_module("dpu-operator_b200")
st = torch.cuda.current_stream().cuda_stream
M = 1 << 20
NEW = os.path.join(ROOT, "dpu-operator_b200", "libb200sfc_new.so")
SECOND = os.environ.get("CFG4_PROBE_SECOND") == "1"
print(json.dumps(dict(library=os.environ.get("B200SFC_LIB", "libb200sfc.so"))), flush=True)
for with_chain in ((True,) if SECOND else (True, False)):
    dp = sfc.Dataplane(0)
    n_ports = 17
    if with_chain:
        import orc
        stg = bench.Stager(dp, orc, sfc)
        n_ports = stg.chain(8, 4)
    spec = sfc.make_gen_spec(**bench.cfg2_spec_kw(1 << 24, n_ports))
    for f in range(0, 1 << 24, 1 << 20):
        dp.flow_add_bulk(*sfc.gen_flows(spec, f, 1 << 20))
    dp.commit()
    for lg in ((26,) if SECOND else (24, 26, 28)):
        n = 1 << lg
        hdr = torch.empty((n, 64), dtype=torch.uint8, device="cuda")
        ports = torch.empty(n, dtype=torch.int16, device="cuda")
        res = torch.zeros(n, dtype=torch.int64, device="cuda")
        hp, pp, rp = hdr.data_ptr(), ports.data_ptr(), res.data_ptr()
        for o in range(0, n, 1 << 22):
            dp.gen_packets(spec, o, 1 << 22, hp + o * 64, pp + o * 2, st)
        torch.cuda.synchronize()
        for flags, name in ((sfc.consts.F_OVERLAP, "overlap"), (sfc.consts.F_OVERLAP | sfc.consts.F_COUNTERS, "overlap+counters")):
            for low in (1, 0):
                dp.set_option("low_regs", low)
                def run():
                    for k in range(n // M):
                        dp.classify(hp + k * M * 64, pp + k * M * 2, rp + k * M * 8, M, flags, st)
                run(); torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0 = time.perf_counter()
                e0.record(); run(); e1.record()
                t_submit = time.perf_counter() - t0
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1)
                print(json.dumps(dict(chain=with_chain, packets=f"2^{lg}", flags=name, low_regs=low, us_per_Mi=round(ms * 1e3 / (n // M), 2),
                                      gpps=round(n / ms / 1e6, 2), host_submit_us_per_launch=round(t_submit * 1e6 / (n // M), 2))), flush=True)
        del hdr, ports, res
        torch.cuda.empty_cache()
    del dp

if not SECOND and os.path.exists(NEW):  # the same measurement with a second build of the library, if one was shipped
    import subprocess
    subprocess.run([sys.executable, os.path.abspath(__file__)], env=dict(os.environ, B200SFC_LIB=NEW, CFG4_PROBE_SECOND="1"))
