"""Rows a9 / a10 / f2: the Intel IPU plugin's rule dialect end to end.

  * the mirrored rule builders (dpu-operator_b200/intel_vsp.py) emit, string for string, the stream that
    tests/golden/make_intel_rules.py computed from the fmt.Sprintf calls in the vendored Go source;
  * every string parses against the reference's p4info and maps onto the B200 tables (at most two tables unmapped);
  * (GPU) the imported rules steer synthetic traffic exactly as the oracle says when fed the same entries.
"""
import importlib
import json
import os

import numpy as np
import pytest

from common import ROOT, Scenario, consts, sfc

intel = importlib.import_module("dpu-operator_b200.intel_vsp")
p4rt = importlib.import_module("dpu-operator_b200.p4rt")
G = json.load(open(os.path.join(ROOT, "tests", "golden", "intel_rules.json")))


def _replay():
    v = intel.IntelVsp(G["acc_macs"])
    for i, m in enumerate(G["vf_macs"]):
        assert v.CreateBridgePort(f"host0-{i}", m) == f"bridge_port/host0-{i}"
    nf = G["nf_macs"]
    v.CreateNetworkFunction(nf["ingress"], nf["egress"], nf["ingress_pr"], nf["egress_pr"])
    return v


def test_rule_stream_equals_the_one_read_out_of_the_go_source():
    v = _replay()
    assert [list(r) for r in v.rules] == G["rules_8_host_vfs_then_1_nf"]
    assert len(v.rules) == 8 * 5 + 2 * 6 + 8 * 2 * 2  # 5 per host VF, 6 per NF port pair, 2 per VSI pair
    assert [list(r) for r in intel.add_peer_to_peer_p4_rules(G["vf_macs"])] == G["rules_peer_to_peer"]
    assert len(G["rules_peer_to_peer"]) == 8 * 7  # |VF| * (|VF| - 1)
    assert intel.VSI_TO_VPORT_OFFSET == G["vsi_to_vport_offset"] == 16
    assert intel.extract_vsi_vport("00:08:00:00:03:14") == (8, 24)  # the example of ipuplugin/bridgeport.go:99-100


def test_bridge_port_argument_checks_follow_the_plugin():
    v = intel.IntelVsp(G["acc_macs"][:1])
    with pytest.raises(ValueError, match="vlan must be within 2-4094"):
        v.CreateBridgePort("host0-0", G["vf_macs"][0], vlan=1)
    with pytest.raises(ValueError, match="2nd octed must be > 0"):
        v.CreateBridgePort("host0-0", "00:00:00:00:03:14")
    v.CreateBridgePort("host0-0", G["vf_macs"][0])
    with pytest.raises(RuntimeError):  # no ACC port representor left
        v.CreateBridgePort("host0-1", G["vf_macs"][1])


class _Sink:
    """What Importer.apply needs of a Dataplane, forwarded to a Scenario so that the oracle receives every rule too."""

    def __init__(self, sc):
        self.sc = sc

    def l2_rule_add(self, in_port, out_port, mac):
        return self.sc.l2_add(in_port, out_port, mac)

    def lpm_add(self, af, prefix, length, port):
        return self.sc.lpm_add(af, prefix, length, port)


def _import(sc, with_routes=True):
    v = _replay()
    imp = p4rt.Importer()
    for table, rule in v.rules + intel.add_peer_to_peer_p4_rules(G["vf_macs"]):
        imp.add(table, rule)
    C = "linux_networking_control."
    if with_routes:  # the LPM / next-hop tables of the same pipeline, in the reference's own grammar
        imp.add(C + "nexthop_table", f"user_meta.cmeta.nexthop_id=7,bit16_zeros=0,action={C}set_nexthop_info_dmac(1,57,0x0011,0x22334455)")
        imp.add(C + "ipv4_table", f"ipv4_table_lpm_root=0,ipv4_dst_match=10.0.0.0/8,action={C}ipv4_set_nexthop_id(7)")
        imp.add(C + "ipv6_table", f"ipv6_table_lpm_root=0,ipv6_dst_match=2001:db8::/32,action={C}ipv6_set_nexthop_id(7)")
        imp.add(C + "tx_lag_table", f"user_meta.cmeta.lag_group_id=0/255,hash=3/7,priority=1,action={C}bypass")  # stays unmapped
    done = imp.apply(_Sink(sc), vsi_mac=v.vsi_mac)
    return v, imp, done


def test_every_generated_rule_maps_onto_the_b200_tables():
    sc = Scenario(sfc.Dataplane(-1))
    v, imp, done = _import(sc)
    assert done["unmapped_tables"] == ["tx_lag_table"] and done["unmapped"] == 1  # <= 2 unmapped tables
    # 8 VFs x 2 + NF 2 x 2 loopbacks, 32 VF<->NF, 56 peer to peer -> exact rules (two of them replace an identical match)
    assert done["l2_exact"] == 16 + 4 + 32 + 56 and done["lpm"] == 2
    # tx_acc_vsi: 8 + 4; source_port_to_pr_map: 8 + 2
    assert done["l2_wildcard"] == 12 + 10
    sc.dp.commit()
    from test_image_walk import Image
    img = Image(sc.dp.table_export())
    vf, acc = intel.mac_bytes(G["vf_macs"][0]), intel.mac_bytes(G["acc_macs"][0])
    # host VF 0 (VSI 0x21) towards its port representor (VSI 0x09): the loopback entry names vport 0x09 + 16
    assert (img.l2(vf[1], acc) >> 32) & 0xFFFF == acc[1] + 16 and (img.l2(vf[1], acc) >> 28) & 0xF == consts.STAGE_L2_EXACT
    # anything else from that VF goes to the representor as well (source_port_to_pr_map through tx_source_port)
    assert (img.l2(vf[1], bytes(6)) >> 32) & 0xFFFF == acc[1] + 16 and (img.l2(vf[1], bytes(6)) >> 28) & 0xF == consts.STAGE_L2_WILD
    # VF 0 -> VF 1 is a peer-to-peer entry
    vf1 = intel.mac_bytes(G["vf_macs"][1])
    assert (img.l2(vf[1], vf1) >> 32) & 0xFFFF == vf1[1] + 16


@pytest.mark.gpu
def test_imported_intel_rules_on_the_gpu(torch_cuda):
    torch = torch_cuda
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    _import(sc)
    dp.commit()
    # frames from every VSI in 9..50 (known and unknown ones) towards MACs 00:<0..63>:00:00:03:14; 40 % non-IP, 15 % IPv6
    spec = sfc.make_gen_spec(n_flows=0, n_ports=42, port_base=9, n_macs=64, nonip_permille=400, v6_permille=150,
                             vlan_permille=100, n_pfx4=0, n_pfx6=0)
    n = 1 << 18
    hdr, ports = sfc.gen_packets_host(spec, 0, n)
    d_hdr, d_ports = torch.from_numpy(hdr).cuda(), torch.from_numpy(ports.view(np.int16)).cuda()
    res = torch.zeros(n, dtype=torch.int64, device="cuda")
    dp.classify(d_hdr.data_ptr(), d_ports.data_ptr(), res.data_ptr(), n, 0, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = res.cpu().numpy().view(np.uint64)
    want = sc.orc.classify(hdr, ports, threads=8)
    assert np.array_equal(got, want), f"{int((got != want).sum())} results differ from the oracle"
    stage = sfc.result_fields(got)[3]
    for s in (consts.STAGE_L2_EXACT, consts.STAGE_L2_WILD, consts.STAGE_MISS):
        assert (stage == s).sum() > 1000, f"stage {s} hardly hit"


def test_physical_port_lag_and_primary_network_builders_equal_the_go_source():
    """The three builders that take their VSI from the IMC (programPhyVportP4Rules, AddLAGP4Rules,
    AddRHPrimaryNetworkVportP4Rules): string for string what the fixture generator evaluated out of the Go source, and
    every string is accepted by the p4rt-ctl grammar against the reference's p4info."""
    assert G["consts"] == dict(mirror_profile_id=intel.MIRROR_PROFILE_ID, bridgeId=intel.BRIDGE_ID, phyPort=intel.PHY_PORT)
    sets = {
        "rules_phy_port_0_pr_vsi_10": intel.program_phy_vport_p4_rules(0, 10),
        "rules_lag": intel.add_lag_p4_rules(),
        "rules_rh_primary_d4_vsi_12_d1_vsi_13": intel.add_rh_primary_network_vport_p4_rules(12, 13, G["rh_d1_mac"]),
    }
    for key, mine in sets.items():
        assert [list(r) for r in mine] == G[key], key
        for table, rule in mine:
            e = p4rt.parse_entry(table, rule)
            assert e.table == table and e.action.startswith("linux_networking_control.")
    # the primary-network set maps onto the B200 tables too: its bridge member and MAC entry become an exact L2 rule
    imp = p4rt.Importer()
    for table, rule in sets["rules_rh_primary_d4_vsi_12_d1_vsi_13"] + sets["rules_phy_port_0_pr_vsi_10"]:
        imp.add(table, rule)
    done = imp.apply(sfc.Dataplane(-1))
    assert done["l2_wildcard"] == 3  # tx_acc_vsi x 2 and source_port_to_pr_map
    assert done["unmapped_tables"] == ["mir_prof", "rx_phy_port_to_pr_map", "rx_source_port"]
