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
