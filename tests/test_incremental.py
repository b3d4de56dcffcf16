"""Row f3: flow churn without rebuilding — sfc_commit's incremental path.  CPU part: an offline handle
patches its host image bucket by bucket and the result must be a table in which every live flow is
findable and every deleted one is gone (checked with the image walker); GPU part: churn between
classify launches, bit-exact against the oracle, through sfc_commit_stream."""
import numpy as np
import pytest

from common import Oracle, Scenario, config3_spec, consts, install_flows, sfc
from test_image_walk import Image


def _flow4_probe(img, k):
    return img.flow(k)[0]


FORMATS = [(1, 1), (1, 2), (2, 0)]  # (flow_format, flow_slots): both sector layouts and the indexed table


@pytest.mark.parametrize("fmt,slots", FORMATS)
def test_incremental_commit_patches_the_image_in_place(fmt, slots):
    dp = sfc.Dataplane(-1)
    dp.set_option("flow_format", fmt)
    dp.set_option("flow_slots", slots)
    spec = config3_spec(n_flows=30000, v6_permille=0, n_ports=8)
    keys, acts = sfc.gen_flows(spec, 0, 30000)
    dp.flow_add_bulk(keys[:20000], acts[:20000])
    dp.commit()
    assert (dp.stat("full_commits"), dp.stat("incremental_commits")) == (1, 0)
    size0, e0 = dp.table_size, dp.table_epoch
    # churn: 3000 new flows, 2000 old ones gone — flows only, so no rebuild
    for k in keys[:4000:2]:  # deletes first: a two-slot table is built at 80 % load and has little headroom
        dp.flow_del(k)
    first = dp.flow_add_bulk(keys[20000:23000], acts[20000:23000])
    assert first == 20000
    dp.commit()
    assert (dp.stat("full_commits"), dp.stat("incremental_commits")) == (1, 1)
    # sectors shipped: one slot per edit plus the cuckoo moves (more at 80 % load); the indexed table ships an index
    # sector and a record sector per edit
    assert 5000 <= dp.stat("patched_buckets") <= {(1, 1): 8000, (1, 2): 14000, (2, 0): 10500}[(fmt, slots)]
    assert dp.table_size == size0 and dp.table_epoch == e0 + 1
    img = Image(dp.table_export())
    assert img.f4_cnt == 21000 and img.flow_format == fmt and (fmt == 2 or img.f4_slots == slots)
    for i in list(range(1, 4000, 2)) + list(range(4000, 23000, 37)):  # alive
        r = _flow4_probe(img, keys[i])
        assert r is not None and r & 0x0FFFFFFF == i, i
    for i in range(0, 4000, 2):  # deleted
        assert _flow4_probe(img, keys[i]) is None, i
    # L2 steering rules do not rebuild either: they go into the idle copy of the L2 tables
    p = dp.port_add("x")
    dp.l2_rule_add(p, p)
    dp.commit()
    assert (dp.stat("full_commits"), dp.stat("incremental_commits")) == (1, 2)
    img2 = Image(dp.table_export())
    assert img2.l2_active == 1 - img.l2_active and img2.l2w_cnt == img.l2w_cnt + 1
    assert img2.l2(p, bytes(6)) >> 56 == consts.ACT_FORWARD and img.l2(p, bytes(6)) is None
    # anything else (LPM, ACL, the chain) forces the rebuild path
    dp.lpm_add(4, bytes([10, 0, 0, 0]), 8, p)
    dp.commit()
    assert dp.stat("full_commits") == 2


def test_growth_falls_back_to_a_rebuild():
    dp = sfc.Dataplane(-1)
    spec = config3_spec(n_flows=20000, v6_permille=300, n_ports=8)
    keys, acts = sfc.gen_flows(spec, 0, 20000)
    dp.flow_add_bulk(keys[:1000], acts[:1000])
    dp.commit()
    dp.flow_add_bulk(keys[1000:], acts[1000:])  # 20x the flows: the table has to grow
    dp.commit()
    assert dp.stat("full_commits") == 2 and dp.stat("incremental_commits") == 0
    img = Image(dp.table_export())
    assert img.f4_cnt + img.f6_cnt == 20000


def test_flow_index_survives_heavy_delete_churn():
    dp = sfc.Dataplane(-1)
    spec = config3_spec(n_flows=50000, v6_permille=200, n_ports=8)
    keys, acts = sfc.gen_flows(spec, 0, 50000)
    dp.flow_add_bulk(keys, acts)
    rng = np.random.default_rng(7)
    order = rng.permutation(50000)
    for i in order[:40000]:
        dp.flow_del(keys[i])
    assert dp.stat("flows") == 10000
    for i in order[:2000]:  # really gone: deleting again fails, adding again works
        with pytest.raises(sfc.SfcError):
            dp.flow_del(keys[i])
    for i in order[40000:42000]:  # really there: adding again is a duplicate
        with pytest.raises(sfc.SfcError) as e:
            dp.flow_add(keys[i], acts[i])
        assert e.value.code == -17
    dp.flow_add_bulk(keys[order[:5000]], acts[order[:5000]])
    assert dp.stat("flows") == 15000


@pytest.mark.gpu
@pytest.mark.parametrize("fmt,slots", FORMATS)
def test_flow_churn_between_launches_on_the_gpu(torch_cuda, fmt, slots):
    torch = torch_cuda
    dp = sfc.Dataplane(0)
    dp.set_option("flow_format", fmt)
    dp.set_option("flow_slots", slots)
    sc = Scenario(dp)
    topo = sc.build_chain(8, 2)
    spec = config3_spec(n_flows=120000, n_ports=topo["n_ports"])
    keys, acts = sfc.gen_flows(spec, 0, 120000)
    sc.flows_add(keys[:80000], acts[:80000])
    dp.commit()
    n = 200000
    hdr, ports = sfc.gen_packets_host(spec, 0, n)
    d_hdr = torch.from_numpy(hdr).cuda()
    d_ports = torch.from_numpy(ports.view(np.int16)).cuda()
    res = torch.zeros(n, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    alive = np.ones(120000, bool)
    alive[80000:] = False
    rng = np.random.default_rng(3)
    nxt = 80000
    for rnd in range(6):
        # add 4000 flows, delete 3000 random live ones; ordered on the stream, no device-wide sync
        first = sc.flows_add(keys[nxt:nxt + 4000], acts[nxt:nxt + 4000])
        assert first == nxt
        alive[nxt:nxt + 4000] = True
        nxt += 4000
        for i in rng.choice(np.nonzero(alive)[0], 3000, replace=False):
            dp.flow_del(keys[i])
            sc.orc.flow_del(keys[i])
            alive[i] = False
        dp.commit_stream(st)
        dp.classify(d_hdr.data_ptr(), d_ports.data_ptr(), res.data_ptr(), n, 0, st)
        torch.cuda.synchronize()
        got = res.cpu().numpy().view(np.uint64)
        want = sc.orc.classify(hdr, ports, threads=8)
        bad = int((got != want).sum())
        assert bad == 0, f"round {rnd}: {bad} results differ after incremental commit"
    if slots != 2:
        assert dp.stat("incremental_commits") == 6 and dp.stat("full_commits") == 1
    else:  # built at 80 % load: the net growth of 1000 flows a round forces a rebuild along the way
        assert dp.stat("incremental_commits") + dp.stat("full_commits") == 7 and dp.stat("incremental_commits") >= 4
    hits = (sfc.result_fields(got)[3] == consts.STAGE_FLOW).mean()
    assert 0.3 < hits < 0.9


def _oracle_from_dump(v):
    """An oracle holding exactly the steering rules the VSP mirror has programmed (rule ids = dump order)."""
    dp, orc = v.dataplane, Oracle()
    mac = lambda s: bytes(int(x, 16) for x in s.split(":"))  # noqa: E731
    for rid, line in enumerate(v.flow_dump()):
        f = dict(kv.split("=", 1) for kv in line.split(",") if "=" in kv)
        inp = dp.port_lookup(f["in_port"])
        out = inp if line.endswith("actions=in_port") else dp.port_lookup(line.rsplit("output:", 1)[1])
        orc.l2_add(inp, out, mac(f["dl_dst"]) if "dl_dst" in f else None, int(f.get("priority", consts.PRIO_MAC)), rid)
    return orc


@pytest.mark.gpu
def test_bridge_port_churn_while_another_thread_classifies(torch_cuda):
    """The reference's real churn (marvell/main.go:347-486): with 1 Mi flows resident, one thread adds and deletes
    a bridge port 1000 times through the VSP mirror — one port and three steering rules per RPC, ONE table commit per
    RPC, never a rebuild — while another thread classifies without pause on its own stream.  Every launch must equal
    the oracle of the rule set before or after the RPC in flight (a launch sees one epoch, never a mixture), and a
    CreateBridgePort must stay under a millisecond."""
    import threading
    import time

    torch = torch_cuda
    v = sfc.VspServer(0, "rpm0")
    nf_macs = ["02:00:00:00:aa:01", "02:00:00:00:aa:02"]
    for m, ifc in zip(nf_macs, ("nfdp0", "nfdp1")):
        v.add_nf_device(m, ifc)
    v.Init(True)
    vf_mac = lambda i: bytes([0x00, i, 0x00, 0x00, 0x03, 0x14])  # noqa: E731
    for i in range(4):
        v.CreateBridgePort(f"host0-{i}", vf_mac(i))
    v.CreateNetworkFunction(*nf_macs)
    dp = v.dataplane
    spec = config3_spec(n_flows=1 << 20, n_ports=12, hit_permille=300, nonip_permille=400, n_macs=6)
    for first in range(0, 1 << 20, 1 << 18):
        dp.flow_add_bulk(*sfc.gen_flows(spec, first, 1 << 18))
    dp.commit()
    full0 = dp.stat("full_commits")
    # the two rule sets the churn alternates between
    orc_a = _oracle_from_dump(v)
    v.CreateBridgePort("host0-4", vf_mac(4))
    orc_b = _oracle_from_dump(v)
    v.DeleteBridgePort("host0-4")
    keys, acts = sfc.gen_flows(spec, 0, 1 << 20)
    for o in (orc_a, orc_b):
        o.flow_add(keys, acts, 0)
    n = 1 << 16
    hdr, ports = sfc.gen_packets_host(spec, 0, n)
    # rule ids grow with every add: compare decisions (action, next NF, port, stage), not the id of an L2 rule
    def decisions(r):
        r = np.asarray(r, dtype=np.uint64)
        l2 = ((r >> np.uint64(28)) & np.uint64(0xF) == consts.STAGE_L2_EXACT) | ((r >> np.uint64(28)) & np.uint64(0xF) == consts.STAGE_L2_WILD)
        return np.where(l2, r & ~np.uint64(0x0FFFFFFF), r)
    want_a, want_b = decisions(orc_a.classify(hdr, ports, threads=8)), decisions(orc_b.classify(hdr, ports, threads=8))
    assert (want_a != want_b).sum() > 100  # the port under churn carries traffic

    stop, errors, launches = threading.Event(), [], [0, 0, 0]

    def classify_loop():
        try:
            torch.cuda.set_device(0)
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                d_hdr = torch.from_numpy(hdr).cuda()
                d_ports = torch.from_numpy(ports.view(np.int16)).cuda()
                res = torch.zeros(n, dtype=torch.int64, device="cuda")
                while not stop.is_set():
                    dp.classify(d_hdr.data_ptr(), d_ports.data_ptr(), res.data_ptr(), n, 0, s.cuda_stream)
                    s.synchronize()
                    got = decisions(res.cpu().numpy().view(np.uint64))
                    a, b = bool(np.array_equal(got, want_a)), bool(np.array_equal(got, want_b))
                    launches[0] += 1
                    launches[1] += a
                    launches[2] += b
                    if not (a or b):
                        errors.append(f"launch {launches[0]}: {int((got != want_a).sum())} words differ from epoch A, "
                                      f"{int((got != want_b).sum())} from epoch B")
                        return
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    t = threading.Thread(target=classify_loop)
    t.start()
    create_s = []
    try:
        for it in range(1000):
            t0 = time.perf_counter()
            v.CreateBridgePort("host0-4", vf_mac(4))
            create_s.append(time.perf_counter() - t0)
            v.DeleteBridgePort("host0-4")
            if errors:
                break
    finally:
        stop.set()
        t.join()
    assert not errors, errors[0]
    assert launches[0] > 50 and launches[1] > 0 and launches[2] > 0, launches  # both epochs were seen by the data plane
    assert dp.stat("full_commits") == full0, "a bridge-port RPC rebuilt the image"
    assert dp.stat("incremental_commits") >= 2000
    med = sorted(create_s)[len(create_s) // 2]
    assert med < 1e-3, f"median CreateBridgePort took {med * 1e3:.2f} ms"


def test_indexed_table_reuses_record_units_under_mixed_family_churn():
    """Format 2 with both families: repeated delete / add rounds patch the image in place, every live flow stays
    findable, every deleted one is gone, and released record units (one per IPv4 flow, a pair per IPv6 flow) are
    handed out again instead of growing the array."""
    dp = sfc.Dataplane(-1)
    dp.set_option("flow_format", 2)
    spec = config3_spec(n_flows=40000, v6_permille=350, n_ports=8)
    keys, acts = sfc.gen_flows(spec, 0, 40000)
    dp.flow_add_bulk(keys[:20000], acts[:20000])
    dp.commit()
    used0 = Image(dp.table_export()).frec_used
    rng = np.random.default_rng(11)
    alive = set(range(20000))
    nxt = 20000
    for rnd in range(5):
        gone = rng.choice(sorted(alive), 2000, replace=False)
        for i in gone:
            dp.flow_del(keys[i])
            alive.discard(int(i))
        dp.flow_add_bulk(keys[nxt:nxt + 2000], acts[nxt:nxt + 2000])
        alive.update(range(nxt, nxt + 2000))
        nxt += 2000
        dp.commit()
        img = Image(dp.table_export())
        assert img.f4_cnt + img.f6_cnt == len(alive) == 20000
        for i in list(gone[:300]):
            assert img.flow(keys[i])[0] is None, (rnd, i)
        for i in rng.choice(sorted(alive), 600, replace=False):
            r = img.flow(keys[i])[0]
            assert r is not None and r & 0x0FFFFFFF != 0xFFFFFFF, (rnd, i)
    assert (dp.stat("full_commits"), dp.stat("incremental_commits")) == (1, 5)
    # 10 000 flows were replaced; without reuse the array would have grown by ~13 500 units
    assert img.frec_used <= used0 + 3000, (used0, img.frec_used)


def test_import_refuses_damaged_flow_index_l2_table_and_results():
    import errno
    import struct as st
    sc = Scenario(sfc.Dataplane(-1))
    sc.dp.set_option("flow_format", 2)
    topo = sc.build_chain(4, 1)
    spec = config3_spec(n_flows=3000, v6_permille=300, n_ports=topo["n_ports"])
    install_flows(sc, spec)
    sc.dp.commit()
    good = sc.dp.table_export()
    img = Image(good)
    sfc.Dataplane(-1).table_import(good)

    def refused(mutate):
        bad = good.copy()
        mutate(bad)
        with pytest.raises(sfc.SfcError) as e:
            sfc.Dataplane(-1).table_import(bad)
        assert e.value.code == -errno.EINVAL
        return str(e.value)

    def put32(buf, off, v):
        buf[off:off + 4] = np.frombuffer(st.pack("<I", v), dtype=np.uint8)

    # an index entry naming a unit beyond the record array
    first = next(o for o in range(img.fidx_off, img.fidx_off + img.fidx_nb * 32, 4) if (o - img.fidx_off) % 32 < 28 and int(img.u32(o, 1)[0]))
    assert "flow index" in refused(lambda b: put32(b, first, (1 << img.fidx_bits) - 1))
    # an L2 exact table with no empty entry left (a miss would probe forever)
    def fill_l2(b):
        for i in range(img.l2x_mask + 1):
            put32(b, img.l2x_off + 16 * i + 12, 0x01FF0001)
    assert "L2" in refused(fill_l2)
    # a stored result with an action the counters cannot index
    hit = next(i for i in range(img.l2x_mask + 1) if int(img.u32(img.l2x_off + 16 * i + 12, 1)[0]))
    assert "unknown action" in refused(lambda b: put32(b, img.l2x_off + 16 * hit + 12, 0x7F000001))
    # a section offset that is no longer aligned for the kernel's 32-byte loads (header field frec_off at byte 424)
    assert "aligned" in refused(lambda b: b.__setitem__(slice(424, 432), np.frombuffer(st.pack("<Q", img.frec_off + 8), dtype=np.uint8)))
