"""BASELINE.json configs[0], replayed: the C++ mirror of the Marvell VSP (csrc/host/), driven through
its C ABI with the object sequence the reference's reconcilers produce from examples/*.yaml, must
program exactly the rule set in tests/golden/replay_rules.json (whose `ovs-ofctl` format strings were
read out of the reference's ovsdp.go — see tests/golden/make_replay.py)."""
import json
import os

import numpy as np
import pytest

from common import ROOT, Oracle, consts, sfc
from packets import SMAC, eth

G = json.load(open(os.path.join(ROOT, "tests", "golden", "replay_rules.json")))


def mac_bytes(s):
    return bytes(int(x, 16) for x in s.split(":"))


def new_vsp(device=-1):
    v = sfc.VspServer(device, G["rpm_port"])
    for m, ifc in zip(G["nf_macs"], G["dp_interfaces"]):
        v.add_nf_device(m, ifc)  # what ConfigureNetworkInterface leaves in deviceStore
    v.Init(dpu_mode=True)
    return v


def bridge_ports(v):
    for i, m in enumerate(G["vf_macs"]):  # HostSideManager.CreateBridgePort(pf=0, vf=i, ...) per workload pod
        assert v.CreateBridgePort(f"host0-{i}", mac_bytes(m)) == f"bridge_port/host0-{i}"


def test_ports_then_chain_gives_the_26_rules():
    v = new_vsp()
    bridge_ports(v)
    assert v.flow_dump() == []  # no NF yet: ports only (main.go:383 `if vsp.isNF`)
    v.CreateNetworkFunction(*G["nf_macs"])  # DpuSideManager pairs the two MACs (dpusidemanager.go:111-116)
    assert v.flow_dump() == G["rules_ports_then_nf"]
    assert len(v.flow_dump()) == 3 * G["num_vfs"] + 2 == 26
    assert v.dataplane.port_list() == G["ports"]


def test_chain_then_ports_gives_the_same_set_in_the_other_order():
    v = new_vsp()
    v.CreateNetworkFunction(*G["nf_macs"])
    bridge_ports(v)
    assert v.flow_dump() == G["rules_nf_then_ports"]
    assert sorted(v.flow_dump()) == sorted(G["rules_ports_then_nf"])


def test_delete_bridge_port_and_network_function():
    v = new_vsp()
    bridge_ports(v)
    v.CreateNetworkFunction(*G["nf_macs"])
    v.DeleteBridgePort(G["deleted_bridge_port"])
    assert v.flow_dump() == G["rules_after_delete_host0_3"]
    assert "vf0-4" not in v.dataplane.port_list()
    v.DeleteNetworkFunction(*G["nf_macs"])  # main.go:575-618: every flow goes, NF ports leave the bridge
    assert v.flow_dump() == []
    assert v.dataplane.port_list() == [p for p in G["ports"] if p not in ("vf0-4", *G["dp_interfaces"])]


def test_error_behaviour_follows_the_reference():
    v = new_vsp()
    with pytest.raises(sfc.SfcError) as e:
        v.CreateBridgePort("eth0", mac_bytes(G["vf_macs"][0]))  # getVFDetails: regexp host(\d+)-(\d+) fails
    assert "no VFId Match Found" in str(e.value)
    w = sfc.VspServer(-1)
    w.Init(dpu_mode=True)
    with pytest.raises(sfc.SfcError):  # unknown NF MAC -> "" interface name -> add-port fails (main.go:493-503)
        w.CreateNetworkFunction("de:ad:be:ef:00:01", "de:ad:be:ef:00:02")
    h = sfc.VspServer(-1)
    h.Init(dpu_mode=False)  # host side programs nothing
    with pytest.raises(sfc.SfcError):
        h.dataplane


def test_replayed_rule_set_forwards_like_the_reachability_matrix():
    """The replayed rules, parsed back from the dumped strings into the oracle, forward every hop of
    the reference's e2e reachability matrix (e2e_test/e2e_test.go:406-513) to the next port."""
    v = new_vsp()
    bridge_ports(v)
    v.CreateNetworkFunction(*G["nf_macs"])
    dp = v.dataplane
    orc = Oracle()
    for rid, line in enumerate(v.flow_dump()):
        f = dict(kv.split("=", 1) for kv in line.split(",") if "=" in kv)
        inp = dp.port_lookup(f["in_port"])
        out = inp if line.endswith("actions=in_port") else dp.port_lookup(line.rsplit("output:", 1)[1])
        prio = int(f.get("priority", consts.PRIO_MAC))
        orc.l2_add(inp, out, mac_bytes(f["dl_dst"]) if "dl_dst" in f else None, prio, rid)
    vf0, inp, out, rpm = (dp.port_lookup(n) for n in ("vf0-1", "dp_interface0", "dp_interface1", "rpm0"))
    m0, ext = mac_bytes(G["vf_macs"][0]), bytes.fromhex("0200000000fe")
    hops = [(ext, vf0, inp), (m0, inp, vf0), (m0, out, out), (ext, out, rpm), (m0, rpm, out)]
    hdr = np.frombuffer(b"".join(eth(d, SMAC, 0x0806, bytes(28)) for d, _, _ in hops), np.uint8).reshape(-1, 64)
    act, _, port, _, _ = sfc.result_fields(orc.classify(hdr, np.array([h[1] for h in hops], np.uint16)))
    assert act.tolist() == [consts.ACT_FORWARD] * len(hops) and port.tolist() == [h[2] for h in hops]
