"""GPU parity: the CUDA path, called through the C ABI, must be bit-exact against the CPU oracle
on the same inputs (integer work: no tolerance)."""
import os

import numpy as np
import pytest

from common import Oracle, Scenario, config2_spec, config3_spec, config4_spec, consts, install_flows, install_lpm_acl, sfc
from packets import handmade_cases

pytestmark = pytest.mark.gpu


def classify_dev(torch, dp, hdr_np, ports_np, flags=0):
    hdr = torch.from_numpy(np.array(hdr_np, dtype=np.uint8, order="C")).cuda()  # a copy: frombuffer views are read-only
    ports = torch.from_numpy(np.array(ports_np, dtype=np.uint16).view(np.int16)).cuda()
    res = torch.zeros(ports_np.size, dtype=torch.int64, device="cuda")
    dp.classify(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), ports_np.size, flags, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return res.cpu().numpy().view(np.uint64)


def gen_dev(torch, dp, spec, first, n):
    hdr = torch.empty((n, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty(n, dtype=torch.int16, device="cuda")
    dp.gen_packets(spec, first, n, hdr.data_ptr(), ports.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return hdr, ports


def assert_same(got, want, hdr=None):
    bad = np.nonzero(got != want)[0]
    if bad.size:
        i = int(bad[0])
        msg = f"{bad.size} of {got.size} results differ; first at packet {i}: got {int(got[i]):#018x} want {int(want[i]):#018x}"
        if hdr is not None:
            msg += f"\nheader: {bytes(np.asarray(hdr)[i]).hex()}"
        raise AssertionError(msg)


def test_generator_host_equals_device(torch_cuda):
    dp = sfc.Dataplane(0)
    spec = config3_spec(n_flows=5000, n_pfx4=100, n_pfx6=50, lpm_permille=500)
    hdr_h, ports_h = sfc.gen_packets_host(spec, 12345, 20000)
    hdr_d, ports_d = gen_dev(torch_cuda, dp, spec, 12345, 20000)
    assert np.array_equal(hdr_d.cpu().numpy(), hdr_h)
    assert np.array_equal(ports_d.cpu().numpy().view(np.uint16), ports_h)


def test_handmade_edge_cases(torch_cuda):
    """Tagged / double tagged / IPv4 options up to the 64-byte limit / fragments / IPv6 / malformed
    version and IHL / non-IP, with rules in every table, against both oracles."""
    dp = sfc.Dataplane(0)
    sc = Scenario(dp, with_py=True)
    hdr, ports = handmade_cases(sc)
    dp.commit()
    got = classify_dev(torch_cuda, dp, hdr, ports)
    want = sc.orc.classify(hdr, ports)
    assert_same(got, want, hdr)
    want_py = np.array([sc.py.classify(bytes(hdr[i]), int(ports[i])) for i in range(len(ports))], dtype=np.uint64)
    assert_same(got, want_py, hdr)
    # every stage must have been exercised by the fixture
    assert set(sfc.result_fields(got)[3].tolist()) == set(range(7))


@pytest.mark.parametrize("n", [1, 31, 127, 128, 129, 4096, 100003])
def test_ragged_sizes(torch_cuda, n):
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    spec = config2_spec()
    install_flows(sc, spec)
    sc.build_chain(8, 1)
    dp.commit()
    hdr, ports = sfc.gen_packets_host(spec, 7, n)
    assert_same(classify_dev(torch_cuda, dp, hdr, ports), sc.orc.classify(hdr, ports), hdr)


def test_empty_batch_and_empty_tables(torch_cuda):
    dp = sfc.Dataplane(0)
    dp.commit()
    dp.classify(0, 0, 0, 0)  # n == 0: no launch, no error
    hdr, ports = sfc.gen_packets_host(config2_spec(), 0, 1000)
    got = classify_dev(torch_cuda, dp, hdr, ports)
    assert np.all(got == np.uint64(consts.RESULT_MISS))


def test_config2_full(torch_cuda):
    """BASELINE.json configs[1]: 1K exact 5-tuple rules, 1M-packet batch."""
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    spec = config2_spec()
    install_flows(sc, spec)
    dp.commit()
    n = 1 << 20
    hdr, ports = gen_dev(torch_cuda, dp, spec, 0, n)
    res = torch_cuda.zeros(n, dtype=torch_cuda.int64, device="cuda")
    dp.classify(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), n, 0, torch_cuda.cuda.current_stream().cuda_stream)
    torch_cuda.cuda.synchronize()
    got = res.cpu().numpy().view(np.uint64)
    want = sc.orc.classify(hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), threads=8)
    assert_same(got, want)
    stage = sfc.result_fields(got)[3]
    hits = int((stage == consts.STAGE_FLOW).sum())
    assert 0.85 * n < hits < 0.95 * n  # 90% of the traffic replays the rule set


def test_config3_batches(torch_cuda):
    """configs[2]: 1M-flow cuckoo table (IPv4 + IPv6), 4096-packet batches, 4-NF chain."""
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    topo = sc.build_chain(8, 4)
    spec = config3_spec(n_flows=1 << 20, n_ports=topo["n_ports"])
    install_flows(sc, spec)
    dp.commit()
    n_batches, b = 64, 4096
    hdr, ports = gen_dev(torch_cuda, dp, spec, 0, n_batches * b)
    res = torch_cuda.zeros(n_batches * b, dtype=torch_cuda.int64, device="cuda")
    st = torch_cuda.cuda.current_stream().cuda_stream
    for i in range(n_batches):  # one launch per batch
        dp.classify(hdr.data_ptr() + i * b * 64, ports.data_ptr() + i * b * 2, res.data_ptr() + i * b * 8, b, 0, st)
    torch_cuda.cuda.synchronize()
    got = res.cpu().numpy().view(np.uint64)
    want = sc.orc.classify(hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), threads=8)
    assert_same(got, want)
    stage = sfc.result_fields(got)[3]
    for s in (consts.STAGE_FLOW, consts.STAGE_L2_EXACT, consts.STAGE_L2_WILD):
        assert (stage == s).sum() > 0, f"stage {s} never hit"


def test_config4_lpm_acl(torch_cuda):
    """configs[3]: LPM (IPv4/IPv6) + ACL on 100K prefixes, mixed traffic."""
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    spec = config4_spec()
    install_lpm_acl(sc, spec)
    dp.commit()
    n = 1 << 22  # the oracle's ACL is a linear scan of 1024 rules per packet: all host threads
    hdr, ports = gen_dev(torch_cuda, dp, spec, 0, n)
    got = classify_dev(torch_cuda, dp, hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16))
    want = sc.orc.classify(hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), threads=os.cpu_count() or 8)
    assert_same(got, want, hdr.cpu().numpy())
    stage = sfc.result_fields(got)[3]
    for s in (consts.STAGE_ACL, consts.STAGE_LPM4, consts.STAGE_LPM6, consts.STAGE_MISS):
        assert (stage == s).sum() > 0, f"stage {s} never hit"


def test_everything_at_once_with_counters(torch_cuda):
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    topo = sc.build_chain(8, 2)
    spec = config3_spec(n_flows=20000, n_ports=topo["n_ports"] + 4, n_pfx4=3000, n_pfx6=1000, lpm_permille=700,
                        n_acl=300, hit_permille=600)
    install_flows(sc, spec)
    install_lpm_acl(sc, spec)
    dp.commit()
    n = 200000
    hdr, ports = sfc.gen_packets_host(spec, 99, n)
    got = classify_dev(torch_cuda, dp, hdr, ports, consts.F_COUNTERS)
    want, cnt = sc.orc.classify(hdr, ports, counters=True)
    assert_same(got, want, hdr)
    assert np.array_equal(dp.counters_read(), cnt)
    assert set(sfc.result_fields(got)[3].tolist()) == set(range(7))
    dp.counters_reset()
    assert dp.counters_read().sum() == 0


def test_classify_host_matches_device_path(torch_cuda):
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    spec = config2_spec()
    install_flows(sc, spec)
    sc.build_chain(8, 1)
    dp.commit()
    n = (1 << 18) * 2 + 4321  # more than two staging chunks, ragged tail
    hdr, ports = sfc.gen_packets_host(spec, 5, n)
    got = dp.classify_host(hdr, ports)
    assert_same(got, sc.orc.classify(hdr, ports, threads=8), hdr)
    assert dp.last_kernel_ms == 0  # kernel timing is opt-in
    dp.set_option("host_timing", 1)
    for bufs, chunk in ((3, 1 << 16), (4, 10000), (2, 1 << 18)):  # other shapes of the staging pipeline
        dp.set_option("host_bufs", bufs)
        dp.set_option("host_chunk", chunk)
        assert np.array_equal(dp.classify_host(hdr, ports), got)
        assert dp.last_kernel_ms > 0


def test_rule_updates_between_commits(torch_cuda):
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    topo = sc.build_chain(4, 1)
    spec = config2_spec(n_flows=2000, n_ports=topo["n_ports"])
    install_flows(sc, spec)
    dp.commit()
    hdr, ports = sfc.gen_packets_host(spec, 0, 50000)
    assert_same(classify_dev(torch_cuda, dp, hdr, ports), sc.orc.classify(hdr, ports), hdr)
    # delete a VF the way DeleteBridgePort does (marvell/main.go:425-486), drop some flows, re-commit
    vf = topo["vfs"][1]
    sc.l2_del(vf, None)
    sc.l2_del(topo["nf_out"][0], sfc.gen_mac(1))
    sc.l2_del(topo["nf_in"][0], sfc.gen_mac(1))
    keys, _ = sfc.gen_flows(spec, 0, 500)
    for k in keys[::3]:
        dp.flow_del(k)
        sc.orc.flow_del(k)
    e0 = dp.table_epoch
    dp.commit()
    assert dp.table_epoch == e0 + 1
    assert_same(classify_dev(torch_cuda, dp, hdr, ports), sc.orc.classify(hdr, ports), hdr)


def test_image_export_import_roundtrip(torch_cuda):
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    topo = sc.build_chain(8, 2)
    spec = config3_spec(n_flows=30000, n_ports=topo["n_ports"])
    install_flows(sc, spec)
    dp.commit()
    hdr, ports = sfc.gen_packets_host(spec, 0, 60000)
    want = classify_dev(torch_cuda, dp, hdr, ports)
    img = dp.table_export()
    dp2 = sfc.Dataplane(0)
    dp2.table_import(img)
    assert_same(classify_dev(torch_cuda, dp2, hdr, ports), want)
    # device-to-device path (what the NCCL broadcast uses)
    buf = torch_cuda.empty(dp.table_size, dtype=torch_cuda.uint8, device="cuda")
    dp.table_export_device(buf.data_ptr(), buf.numel())
    torch_cuda.cuda.synchronize()
    dp3 = sfc.Dataplane(0)
    dp3.table_import_device(buf.data_ptr(), buf.numel())
    assert_same(classify_dev(torch_cuda, dp3, hdr, ports), want)
    with pytest.raises(sfc.SfcError):
        dp2.table_import(img[:-256])


def test_flow_roundtrip_property_16m(torch_cuda):
    """Full-size property (configs[4] table): with 16M flows resident, every packet that replays
    flow f must come back with rule index f and that flow's action — checked without the oracle by
    regenerating flow f from the index the GPU returned."""
    dp = sfc.Dataplane(0)
    spec = config3_spec(n_flows=1 << 24, n_ports=8, nonip_permille=0, hit_permille=1000, zipf=0)
    chunk = 1 << 20
    for first in range(0, 1 << 24, chunk):
        keys, acts = sfc.gen_flows(spec, first, chunk)
        assert dp.flow_add_bulk(keys, acts) == first
    dp.commit()
    n = 1 << 22
    hdr, ports = gen_dev(torch_cuda, dp, spec, 0, n)
    res = torch_cuda.zeros(n, dtype=torch_cuda.int64, device="cuda")
    dp.classify(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), n, 0, torch_cuda.cuda.current_stream().cuda_stream)
    torch_cuda.cuda.synchronize()
    got = res.cpu().numpy().view(np.uint64)
    act, nf, port, stage, idx = sfc.result_fields(got)
    hdr_h = hdr.cpu().numpy()
    # fragments (1/256 of IPv4 packets) lose their ports by design and may miss; everything else hits
    hit = stage == consts.STAGE_FLOW
    assert hit.mean() > 0.99
    sample = np.nonzero(hit)[0][:: max(1, hit.sum() // 200000)]
    for i in sample[:20000]:
        keys, acts = sfc.gen_flows(spec, int(idx[i]), 1)
        p = Oracle.parse(hdr_h[i])
        k = keys[0]
        w = 4 if k["ip_ver"] == 4 else 16
        assert (p["ver"], p["vlan"], p["proto"], p["sport"], p["dport"]) == (k["ip_ver"], k["vlan"], k["proto"], k["sport"], k["dport"])
        assert p["src"][:w] == bytes(k["src"][:w]) and p["dst"][:w] == bytes(k["dst"][:w])
        assert act[i] == acts[0]["action"] and port[i] == acts[0]["out_port"]


def test_repeated_launches_are_stable(torch_cuda):
    """Race guard.  An early version released a shared-memory tile to the TMA producer before the
    consumers' LDS data had actually returned; it showed up as a few wrong rows once in hundreds of
    launches.  Hammer the kernel with back-to-back launches over rotating batches and demand
    bit-exact results every time."""
    torch = torch_cuda
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    spec = config2_spec()
    install_flows(sc, spec)
    dp.commit()
    pool, n = 4, 1 << 20
    hdr = torch.empty((pool, n, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((pool, n), dtype=torch.int16, device="cuda")
    res = torch.zeros((pool, n), dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for b in range(pool):
        dp.gen_packets(spec, b * n, n, hdr[b].data_ptr(), ports[b].data_ptr(), st)
    torch.cuda.synchronize()
    want = torch.stack([
        torch.from_numpy(sc.orc.classify(hdr[b].cpu().numpy(), ports[b].cpu().numpy().view(np.uint16), threads=8).view(np.int64))
        for b in range(pool)]).cuda()
    for it in range(25):
        res.zero_()
        for i in range(pool * 8):
            b = i % pool
            dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), n, 0, st)
        torch.cuda.synchronize()
        bad = int((res != want).sum().item())
        assert bad == 0, f"iteration {it}: {bad} results differ from the oracle"


def test_vsp_replay_on_the_gpu(torch_cuda):
    """configs[0] end to end: the rules programmed through the VSP mirror (CreateBridgePort /
    CreateNetworkFunction, the reference's own RPCs) steer synthetic traffic on the GPU exactly as
    the oracle says when fed the dumped rule strings."""
    import json
    import os

    from common import ROOT

    G = json.load(open(os.path.join(ROOT, "tests", "golden", "replay_rules.json")))
    mac = lambda s: bytes(int(x, 16) for x in s.split(":"))  # noqa: E731
    v = sfc.VspServer(0, G["rpm_port"])
    for m, ifc in zip(G["nf_macs"], G["dp_interfaces"]):
        v.add_nf_device(m, ifc)
    v.Init(True)
    for i, m in enumerate(G["vf_macs"][:-1]):
        v.CreateBridgePort(f"host0-{i}", mac(m))
    v.CreateNetworkFunction(*G["nf_macs"])
    v.CreateBridgePort(f"host0-{len(G['vf_macs']) - 1}", mac(G["vf_macs"][-1]))  # one pod arrives after the chain
    assert sorted(v.flow_dump()) == sorted(G["rules_ports_then_nf"])
    dp = v.dataplane
    orc = Oracle()
    for rid, line in enumerate(v.flow_dump()):
        f = dict(kv.split("=", 1) for kv in line.split(",") if "=" in kv)
        inp = dp.port_lookup(f["in_port"])
        out = inp if line.endswith("actions=in_port") else dp.port_lookup(line.rsplit("output:", 1)[1])
        orc.l2_add(inp, out, mac(f["dl_dst"]) if "dl_dst" in f else None, int(f.get("priority", consts.PRIO_MAC)), rid)
    spec = sfc.make_gen_spec(n_flows=0, n_ports=len(G["ports"]) + 2, n_macs=10, nonip_permille=300, v6_permille=200,
                             vlan_permille=200)
    hdr, ports = sfc.gen_packets_host(spec, 0, 100000)
    got = classify_dev(torch_cuda, dp, hdr, ports)
    want = orc.classify(hdr, ports)
    # rule ids differ (the library numbers rules in add order, the dump is in list order): compare decisions
    ga, gn, gp, gs, _ = sfc.result_fields(got)
    wa, wn, wp, ws, _ = sfc.result_fields(want)
    assert np.array_equal(ga, wa) and np.array_equal(gp, wp) and np.array_equal(gs, ws)
    assert (gs == consts.STAGE_L2_EXACT).sum() > 0 and (gs == consts.STAGE_L2_WILD).sum() > 0 and (gs == consts.STAGE_MISS).sum() > 0


def test_grpc_vsp_drives_the_gpu_dataplane(torch_cuda, tmp_path):
    """Row f1 end to end: rules arrive over the reference's gRPC plugin boundary and packets are then
    classified on the GPU by the dataplane the VSP owns."""
    import importlib
    import json
    import os

    from common import ROOT
    from packets import SMAC, eth

    vs = importlib.import_module("dpu-operator_b200.vsp_server")
    G = json.load(open(os.path.join(ROOT, "tests", "golden", "replay_rules.json")))
    mac = lambda s: bytes(int(x, 16) for x in s.split(":"))  # noqa: E731
    addr = f"unix:{tmp_path}/vsp.sock"
    server, impl = vs.serve(addr, device=0, rpm_port=G["rpm_port"])
    try:
        c = vs.VspClient(addr)
        c.Init(vs.msg("Vendor.InitRequest")(dpu_mode=True))
        for i, m in enumerate(G["vf_macs"]):
            req = vs.msg(f"{vs.OPI}.CreateBridgePortRequest")()
            req.bridge_port.name, req.bridge_port.spec.mac_address = f"host0-{i}", mac(m)
            c.CreateBridgePort(req)
        c.CreateNetworkFunction(vs.msg("Vendor.NFRequest")(input=G["nf_macs"][0], output=G["nf_macs"][1]))
        dp = impl.vsp.dataplane
        vf0, inp, out, rpm = (dp.port_lookup(n) for n in ("vf0-1", "dp_interface0", "dp_interface1", "rpm0"))
        m0, ext = mac(G["vf_macs"][0]), bytes.fromhex("0200000000fe")
        hops = [(ext, vf0, inp), (m0, inp, vf0), (m0, out, out), (ext, out, rpm), (m0, rpm, out), (ext, inp, None)]
        hdr = np.frombuffer(b"".join(eth(d, SMAC, 0x0806, bytes(28)) for d, _, _ in hops), np.uint8).reshape(-1, 64)
        got = classify_dev(torch_cuda, dp, hdr, np.array([h[1] for h in hops], np.uint16))
        act, _, port, _, _ = sfc.result_fields(got)
        assert act.tolist() == [consts.ACT_FORWARD] * 5 + [consts.ACT_TO_CP]
        assert port.tolist()[:5] == [h[2] for h in hops[:5]]
        c.close()
    finally:
        server.stop(0)


def test_three_hash_dense_table(torch_cuda):
    """flow_hashes = 3 at 80 % load: the third bucket choice and the denser table must stay exact."""
    dp = sfc.Dataplane(0)
    dp.set_option("flow_hashes", 3)
    dp.set_option("flow_load_pct", 80)
    sc = Scenario(dp)
    spec = config3_spec(n_flows=100000, n_ports=8, v6_permille=300)
    install_flows(sc, spec)
    dp.commit()
    hdr, ports = sfc.gen_packets_host(spec, 0, 150000)
    assert_same(classify_dev(torch_cuda, dp, hdr, ports), sc.orc.classify(hdr, ports, threads=8), hdr)


@pytest.mark.parametrize("feat", [None, "1", "7"])
def test_acl_3000_rules_single_tree_and_forests(torch_cuda, feat, monkeypatch):
    """A larger ACL, once with the builder's own choice and once with the rule set forced into
    category forests (SFC_ACL_FEAT): every shape of the decision structure must give the oracle's verdict."""
    if feat is not None:
        monkeypatch.setenv("SFC_ACL_FEAT", feat)
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    spec = config4_spec(n_pfx4=2000, n_pfx6=500, n_acl=3000)
    install_lpm_acl(sc, spec)
    dp.commit()
    hdr, ports = sfc.gen_packets_host(spec, 0, 60000)
    got = classify_dev(torch_cuda, dp, hdr, ports, consts.F_COUNTERS)
    want, cnt = sc.orc.classify(hdr, ports, counters=True)
    assert_same(got, want, hdr)
    assert np.array_equal(dp.counters_read(), cnt)
    assert (sfc.result_fields(got)[3] == consts.STAGE_ACL).sum() > 100


def test_two_threads_two_streams_share_one_handle(torch_cuda):
    """Data-plane threads, one stream each, may share a handle (header: one caller thread per stream)."""
    import threading
    t = torch_cuda
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    spec = config2_spec()
    install_flows(sc, spec)
    dp.commit()
    n, rounds = 1 << 16, 200
    hdr, ports = gen_dev(t, dp, spec, 0, 2 * n)
    want = sc.orc.classify(hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), threads=8)
    res = t.zeros((2, rounds, n), dtype=t.int64, device="cuda")
    streams = [t.cuda.Stream(), t.cuda.Stream()]
    errors = []

    def worker(i):
        try:
            for r in range(rounds):
                dp.classify(hdr.data_ptr() + i * n * 64, ports.data_ptr() + i * n * 2, res[i, r].data_ptr(), n,
                            consts.F_OVERLAP if r % 2 else 0, streams[i].cuda_stream)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(2)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    t.cuda.synchronize()
    assert not errors, errors
    for i in range(2):
        w = t.from_numpy(want[i * n:(i + 1) * n].view(np.int64).copy()).cuda()
        assert bool((res[i] == w).all()), f"thread {i}: some launch produced different results"


def test_two_slot_dense_tables(torch_cuda):
    """Option flow_slots = 2: dense two-slot buckets for both families; same answers, smaller table."""
    sizes = {}
    for slots in (1, 2):
        dp = sfc.Dataplane(0)
        dp.set_option("flow_slots", slots)
        sc = Scenario(dp)
        topo = sc.build_chain(8, 3)
        spec = config3_spec(n_flows=300000, n_ports=topo["n_ports"])
        install_flows(sc, spec)
        dp.commit()
        sizes[slots] = dp.table_size
        hdr, ports = gen_dev(torch_cuda, dp, spec, 0, 1 << 19)
        got = classify_dev(torch_cuda, dp, hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), consts.F_COUNTERS)
        want, cnt = sc.orc.classify(hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), counters=True)
        assert_same(got, want, hdr.cpu().numpy())
        assert np.array_equal(dp.counters_read(), cnt)
        assert (sfc.result_fields(got)[3] == consts.STAGE_FLOW).mean() > 0.5
    assert sizes[2] < 0.6 * sizes[1]


def test_host_submit_wait_pipeline(torch_cuda):
    """sfc_classify_host_submit / _wait: several batches in flight, collected in order, equal to the blocking call."""
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    spec = config2_spec()
    install_flows(sc, spec)
    dp.commit()
    sizes = [300001, 1 << 18, 77, 600000, 1]
    batches = [sfc.gen_packets_host(spec, 1000 * i, n) for i, n in enumerate(sizes)]
    want = [dp.classify_host(h, p) for h, p in batches]
    for bufs in (2, 4):
        dp.set_option("host_bufs", bufs)
        hs = [np.ascontiguousarray(h) for h, _ in batches]
        ps = [np.ascontiguousarray(p) for _, p in batches]
        outs = [np.zeros(n, dtype=np.uint64) for n in sizes]
        tickets = [dp.classify_host_submit(h.ctypes.data, p.ctypes.data, o.ctypes.data, o.size) for h, p, o in zip(hs, ps, outs)]
        assert tickets == sorted(tickets)
        dp.classify_host_wait(tickets[1])  # the first two calls are complete, later ones may not be
        assert np.array_equal(outs[0], want[0]) and np.array_equal(outs[1], want[1])
        dp.classify_host_wait()  # everything
        for o, w in zip(outs, want):
            assert np.array_equal(o, w)
    assert_same(want[0], sc.orc.classify(*batches[0], threads=8))


def test_launches_survive_cuda_graph_capture_and_replay(torch_cuda):
    """A poll loop captured once and replayed: sfc_classify and sfc_classify_batches allocate nothing and their
    ticket slots reset themselves, so a captured sequence can be replayed any number of times."""
    t = torch_cuda
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    topo = sc.build_chain(8, 2)
    spec = config3_spec(n_flows=50000, n_ports=topo["n_ports"])
    install_flows(sc, spec)
    dp.commit()
    nb, b = 24, 4096
    hdr, ports = gen_dev(t, dp, spec, 0, nb * b)
    want = t.from_numpy(sc.orc.classify(hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), threads=8).view(np.int64).copy()).cuda()
    res = t.zeros(nb * b, dtype=t.int64, device="cuda")
    # warm up outside capture (first launches set function attributes)
    st = t.cuda.current_stream().cuda_stream
    dp.classify(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), b, consts.F_OVERLAP, st)
    dp.classify_batches([(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), b)], consts.F_OVERLAP, st)
    t.cuda.synchronize()
    g = t.cuda.CUDAGraph()
    with t.cuda.graph(g):
        s = t.cuda.current_stream().cuda_stream
        for i in range(nb // 2):  # first half: one launch per batch
            dp.classify(hdr.data_ptr() + i * b * 64, ports.data_ptr() + i * b * 2, res.data_ptr() + i * b * 8, b, consts.F_OVERLAP, s)
        rest = [(hdr.data_ptr() + i * b * 64, ports.data_ptr() + i * b * 2, res.data_ptr() + i * b * 8, b) for i in range(nb // 2, nb)]
        dp.classify_batches(rest, consts.F_OVERLAP, s)  # second half: one launch for all
    for _ in range(5):
        res.zero_()
        g.replay()
        t.cuda.synchronize()
        assert t.equal(res, want)


def test_indexed_flow_format_on_the_gpu(torch_cuda):
    """Option flow_format = 2 (fingerprint index + dense records, one code path for IPv4 and IPv6): same answers and
    counters as the oracle on mixed traffic, in an image about half the size."""
    sizes = {}
    for fmt in (1, 2):
        dp = sfc.Dataplane(0)
        dp.set_option("flow_format", fmt)
        sc = Scenario(dp)
        topo = sc.build_chain(8, 3)
        spec = config3_spec(n_flows=300000, n_ports=topo["n_ports"])
        install_flows(sc, spec)
        dp.commit()
        sizes[fmt] = dp.table_size
        hdr, ports = gen_dev(torch_cuda, dp, spec, 0, 1 << 19)
        got = classify_dev(torch_cuda, dp, hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), consts.F_COUNTERS)
        want, cnt = sc.orc.classify(hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), threads=8, counters=True)
        assert_same(got, want, hdr.cpu().numpy())
        assert np.array_equal(dp.counters_read(), cnt)
        assert (sfc.result_fields(got)[3] == consts.STAGE_FLOW).mean() > 0.5
    assert sizes[2] < 0.65 * sizes[1]


def test_config5_16m_flows_against_the_oracle(torch_cuda):
    """configs[4] table at full size — 16 Mi flows (20 % IPv6) resident, Zipf traffic, 4-NF chain, counters on —
    compared word for word with the oracle holding the same 16 Mi flows, on 2^22 packets; the flows go into the
    oracle from ITS generator (oracle/workload.cpp), into the library from the product's."""
    from common import OracleGen
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    topo = sc.build_chain(8, 4)
    nf = 1 << 24
    spec = config3_spec(n_flows=nf, n_ports=topo["n_ports"])
    for first in range(0, nf, 1 << 20):
        assert dp.flow_add_bulk(*sfc.gen_flows(spec, first, 1 << 20)) == first
        sc.orc.flow_add(*OracleGen.flows(spec, first, 1 << 20), first)
    dp.commit()
    n = 1 << 22
    hdr, ports = gen_dev(torch_cuda, dp, spec, 5 << 22, n)
    got = classify_dev(torch_cuda, dp, hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), consts.F_COUNTERS)
    oh, op = OracleGen.packets(spec, 5 << 22, n)
    want, cnt = sc.orc.classify(oh, op, threads=os.cpu_count() or 8, counters=True)
    assert_same(got, want, oh)
    assert np.array_equal(dp.counters_read(), cnt)
    stage = sfc.result_fields(got)[3]
    assert (stage == consts.STAGE_FLOW).mean() > 0.8 and (stage == consts.STAGE_MISS).sum() > 0


@pytest.mark.parametrize("with_acl,low_regs", [(True, -1), (False, 0), (False, 1)])
def test_every_kernel_variant_is_stable_under_overlapped_launches(torch_cuda, with_acl, low_regs):
    """Race guard for the variants test_repeated_launches_are_stable does not reach: the one-stage ACL kernels, the
    40-register kernels, the multi-batch kernels and the counters variants, all under overlapped (programmatic
    dependent) launches over rotating buffers; every launch's full result array is compared with the oracle's."""
    torch = torch_cuda
    F = consts
    n, pool, iters = 1 << 18, 6, 4
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    dp.set_option("low_regs", low_regs)
    topo = sc.build_chain(8, 3)
    spec = config3_spec(n_flows=200000, n_ports=topo["n_ports"] + 2, n_pfx4=5000, n_pfx6=2000, lpm_permille=600,
                        n_acl=200 if with_acl else 0, hit_permille=700)
    install_flows(sc, spec)
    install_lpm_acl(sc, spec)
    dp.commit()
    st = torch.cuda.current_stream().cuda_stream
    hdr = torch.empty((pool, n, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((pool, n), dtype=torch.int16, device="cuda")
    res = torch.zeros((pool, n), dtype=torch.int64, device="cuda")
    for b in range(pool):
        dp.gen_packets(spec, b * n, n, hdr[b].data_ptr(), ports[b].data_ptr(), st)
    torch.cuda.synchronize()
    want = torch.stack([torch.from_numpy(sc.orc.classify(hdr[b].cpu().numpy(), ports[b].cpu().numpy().view(np.uint16),
                                                         threads=os.cpu_count() or 8).view(np.int64)) for b in range(pool)]).cuda()
    cuts = [0, 4096, 4096 + 255, 70000, 70000 + 513, 200000, n]
    for mode in ("classify", "classify+counters", "batches", "batches+counters"):
        flags = F.F_OVERLAP | (F.F_COUNTERS if "counters" in mode else 0)
        for it in range(iters):
            res.zero_()
            for i in range(pool * 6):
                b = i % pool
                if mode.startswith("batches"):
                    dp.classify_batches([(hdr[b].data_ptr() + a * 64, ports[b].data_ptr() + a * 2, res[b].data_ptr() + a * 8, e - a)
                                         for a, e in zip(cuts, cuts[1:])], flags, st)
                else:
                    dp.classify(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr(), n, flags, st)
            torch.cuda.synchronize()
            bad = int((res != want).sum())
            assert bad == 0, f"{mode}, iteration {it}: {bad} results differ from the oracle"
