"""The rule builders of the Intel IPU vendor plugin, mirrored: which P4Runtime table entries the plugin programs
for a host VF, for a network function and between peers, as the rule strings it hands to p4rt-ctl.

Reference: vendored github.com/intel/ipu-opi-plugins/ipu-plugin, pkg/p4rtclient/p4rtclient.go
  AddHostVfP4Rules           :647-731   five entries per host VF (called by CreateBridgePort, ipuplugin/bridgeport.go:151)
  programNfPrVportP4Rules    :272-340   six entries per NF port / port-representor pair
  programVsiToVsiP4Rules     :412-444   two loopback entries per VSI pair
  AddNFP4Rules               :819-859   2 x programNfPrVportP4Rules, then every VF x {ingress, egress}
  AddPeerToPeerP4Rules       :903-919   every unordered pair of host VFs
  programPhyVportP4Rules     :123-181   a physical port and the port representor that fronts it
  AddLAGP4Rules              :939-976   the LPM root look-up entry and the eight LAG hash buckets
  AddRHPrimaryNetworkVportP4Rules :1017-1089  mirror profile, primary-network bridge and MAC entry
(the last three look a VSI up on the IMC over ssh, `ImcQueryfindVsiGivenMacAddr`; here the VSI is an argument)
and pkg/utils/utils.go :39 (vsiToVportOffset), :76-85 (ExtractVsiVportInfo: VSI = second MAC octet), :105-107.

The reference is Go and executes `p4rt-ctl add-entry br0 <table> "<rule>"` once per entry; here the builders
return the (table, rule) pairs, and `dpu-operator_b200.p4rt.Importer` turns them into table updates of the B200
dataplane.  Python, not C++, because what consumes the strings — the reference's p4rt-ctl — is Python too, and
the parser that stands in for it here (p4rt.py) is pinned to it byte for byte.  tests/test_intel_vsp.py holds
every string to tests/golden/intel_rules.json, which tests/golden/make_intel_rules.py produces by reading the
fmt.Sprintf calls out of the Go source.
"""
from __future__ import annotations

VSI_TO_VPORT_OFFSET = 16  # utils.go:39
MIRROR_PROFILE_ID, BRIDGE_ID, PHY_PORT = 3, 0, 0  # p4rtclient.go:30-33
_C = "linux_networking_control."


def mac_bytes(mac) -> bytes:
    if isinstance(mac, (bytes, bytearray)):
        b = bytes(mac)
    else:
        b = bytes(int(x, 16) for x in str(mac).split(":"))
    if len(b) != 6:
        raise ValueError(f"invalid mac address {mac!r}")
    return b


def extract_vsi_vport(mac) -> tuple[int, int]:
    """utils.ExtractVsiVportInfo (:76-85): the VSI is the MAC's second octet, the vport is VSI + 16 (:105-107)."""
    vsi = mac_bytes(mac)[1]
    return vsi, vsi + VSI_TO_VPORT_OFFSET


def add_host_vf_p4_rules(host_vf_mac, acc_mac) -> list[tuple[str, str]]:
    """AddHostVfP4Rules (:647-731): a host VF and the ACC port representor that fronts it on the bridge."""
    vf_vsi, vf_vport = extract_vsi_vport(host_vf_mac)
    pr_vsi, pr_vport = extract_vsi_vport(acc_mac)
    return [
        (_C + "tx_source_port", f"vmeta.common.vsi={vf_vsi}/0x7ff,priority=1,action={_C}set_source_port({vf_vport})"),
        (_C + "tx_acc_vsi", f"vmeta.common.vsi={pr_vsi},zero_padding=0,action={_C}l2_fwd_and_bypass_bridge({vf_vport})"),
        (_C + "vsi_to_vsi_loopback", f"vmeta.common.vsi={pr_vsi},target_vsi={vf_vsi},action={_C}fwd_to_vsi({vf_vport})"),
        (_C + "vsi_to_vsi_loopback", f"vmeta.common.vsi={vf_vsi},target_vsi={pr_vsi},action={_C}fwd_to_vsi({pr_vport})"),
        (_C + "source_port_to_pr_map", f"user_meta.cmeta.source_port={vf_vport},zero_padding=0,action={_C}fwd_to_vsi({pr_vport})"),
    ]


def program_nf_pr_vport_p4_rules(ingress_mac, egress_mac) -> list[tuple[str, str]]:
    """programNfPrVportP4Rules (:272-340): called with (NF port MAC, its port representor's MAC)."""
    in_vsi, in_vport = extract_vsi_vport(ingress_mac)
    eg_vsi, eg_vport = extract_vsi_vport(egress_mac)
    return [
        (_C + "tx_source_port", f"vmeta.common.vsi={in_vsi}/0x7ff,priority=1,action={_C}set_source_port({in_vport})"),
        (_C + "tx_acc_vsi", f"vmeta.common.vsi={in_vsi},zero_padding=0,action={_C}l2_fwd_and_bypass_bridge({eg_vport})"),
        (_C + "tx_acc_vsi", f"vmeta.common.vsi={eg_vsi},zero_padding=0,action={_C}l2_fwd_and_bypass_bridge({in_vport})"),
        (_C + "vsi_to_vsi_loopback", f"vmeta.common.vsi={eg_vsi},target_vsi={in_vsi},action={_C}fwd_to_vsi({in_vport})"),
        (_C + "vsi_to_vsi_loopback", f"vmeta.common.vsi={in_vsi},target_vsi={eg_vsi},action={_C}fwd_to_vsi({eg_vport})"),
        (_C + "source_port_to_pr_map", f"user_meta.cmeta.source_port={in_vport},zero_padding=0,action={_C}fwd_to_vsi({eg_vport})"),
    ]


def program_vsi_to_vsi_p4_rules(mac1, mac2) -> list[tuple[str, str]]:
    """programVsiToVsiP4Rules (:412-444): the two directions between two VSIs."""
    v1, p1 = extract_vsi_vport(mac1)
    v2, p2 = extract_vsi_vport(mac2)
    return [
        (_C + "vsi_to_vsi_loopback", f"vmeta.common.vsi={v1},target_vsi={v2},action={_C}fwd_to_vsi({p2})"),
        (_C + "vsi_to_vsi_loopback", f"vmeta.common.vsi={v2},target_vsi={v1},action={_C}fwd_to_vsi({p1})"),
    ]


def add_nf_p4_rules(vf_macs, ingress_mac, egress_mac, ingress_pr_mac, egress_pr_mac) -> list[tuple[str, str]]:
    """AddNFP4Rules (:819-859), called by CreateNetworkFunction (ipuplugin/networkfunctionservice.go:45-71)."""
    out = program_nf_pr_vport_p4_rules(ingress_mac, ingress_pr_mac) + program_nf_pr_vport_p4_rules(egress_mac, egress_pr_mac)
    for vf in vf_macs:
        for nf in (ingress_mac, egress_mac):
            out += program_vsi_to_vsi_p4_rules(vf, nf)
    return out


def add_peer_to_peer_p4_rules(vf_macs) -> list[tuple[str, str]]:
    """AddPeerToPeerP4Rules (:903-919): every unordered pair of host VFs, |VF| * (|VF| - 1) entries."""
    out = []
    for i in range(len(vf_macs)):
        for j in range(i + 1, len(vf_macs)):
            out += program_vsi_to_vsi_p4_rules(vf_macs[i], vf_macs[j])
    return out


def program_phy_vport_p4_rules(phy_port: int, pr_vsi: int) -> list[tuple[str, str]]:
    """programPhyVportP4Rules (:123-181); pr_vsi is what the IMC reports for the port representor's MAC."""
    pr_vport = pr_vsi + VSI_TO_VPORT_OFFSET
    return [
        (_C + "rx_source_port", f"vmeta.common.port_id={phy_port},zero_padding=0,action={_C}set_source_port({phy_port})"),
        (_C + "source_port_to_pr_map", f"user_meta.cmeta.source_port={phy_port},zero_padding=0,action={_C}fwd_to_vsi({pr_vport})"),
        (_C + "tx_acc_vsi", f"vmeta.common.vsi={pr_vsi},zero_padding=0,action={_C}l2_fwd_and_bypass_bridge({phy_port})"),
    ]


def add_lag_p4_rules() -> list[tuple[str, str]]:
    """AddLAGP4Rules (:939-976): 1 + 8 entries."""
    out = [(_C + "ipv4_lpm_root_lut", f"user_meta.cmeta.bit16_zeros=4/65535,priority=2048,action={_C}ipv4_lpm_root_lut_action(0)")]
    for idx in range(8):
        out.append((_C + "tx_lag_table", f"user_meta.cmeta.lag_group_id=0/255,hash={idx}/7,priority=1,action={_C}bypass"))
    return out


def add_rh_primary_network_vport_p4_rules(d4_vsi: int, d1_vsi: int, d1_mac) -> list[tuple[str, str]]:
    """AddRHPrimaryNetworkVportP4Rules (:1017-1089); the two VSIs are what the IMC reports for the D4 and D1 MACs,
    the MAC goes into the rule with its colons stripped (getStrippedMacAndVsi :103-121)."""
    mac = mac_bytes(d1_mac).hex()
    return [
        (_C + "mir_prof", f"mirror_prof_key={MIRROR_PROFILE_ID},action={_C}mir_prof_action(vport_id={d1_vsi},mode=0,port_dest_type=0,"
                          f"dest_id={d1_vsi},func_valid=1,store_vsi=1)"),
        (_C + "rx_phy_port_to_pr_map", f"vmeta.common.port_id=0x00,zero_padding=0x0000,action={_C}mirror_and_send({d4_vsi + 16},{MIRROR_PROFILE_ID})"),
        (_C + "tx_acc_vsi", f"vmeta.common.vsi={d1_vsi},zero_padding=0,action={_C}l2_fwd_and_bypass_bridge({PHY_PORT})"),
        (_C + "source_port_to_bridge_map", f"user_meta.cmeta.source_port={PHY_PORT}/0xffff,hdrs.vlan_ext[vmeta.common.depth].hdr.vid={PHY_PORT}/0xfff,"
                                           f"priority=1,action={_C}set_bridge_id(bridge_id={BRIDGE_ID})"),
        (_C + "l2_fwd_rx_table", f"user_meta.pmeta.bridge_id={BRIDGE_ID},dst_mac=0x{mac},action={_C}l2_fwd({d1_vsi + 16})"),
    ]


class IntelVsp:
    """The call order of the Intel plugin's two RPCs, on top of the builders: CreateBridgePort programs the host VF
    against the next free ACC port representor (bridgeport.go:93-158), CreateNetworkFunction programs the NF ports
    and their loopbacks to every VF seen so far (networkfunctionservice.go:45-71).  `rules` is the stream, in the
    order the plugin would have executed p4rt-ctl; `vsi_mac` is what a consumer needs to turn target_vsi back into a
    destination MAC (the pipeline derives it from the frame; the rule strings no longer carry it)."""

    def __init__(self, acc_macs):
        self.acc_macs = [mac_bytes(m) for m in acc_macs]  # AccApfMacList: MACs of the ACC port representors
        self.next_acc = 0
        self.vf_macs: list[bytes] = []
        self.rules: list[tuple[str, str]] = []
        self.vsi_mac: dict[int, bytes] = {}

    def _seen(self, mac):
        m = mac_bytes(mac)
        self.vsi_mac[m[1]] = m
        return m

    def CreateBridgePort(self, name: str, mac, vlan: int = 2):
        m = self._seen(mac)
        if not 2 <= vlan <= 4094:
            raise ValueError(f"invalid vlan {vlan}, vlan must be within 2-4094 range")  # bridgeport.go:112-115
        if m[1] < 1:
            raise ValueError(f"invalid VSI:{m[1]} in given mac address, the value in 2nd octed must be > 0")  # :117-120
        if self.next_acc >= len(self.acc_macs):
            raise RuntimeError("error from allocateAccInterface")
        acc = self._seen(self.acc_macs[self.next_acc])
        self.next_acc += 1
        self.vf_macs.append(m)
        self.rules += add_host_vf_p4_rules(m, acc)
        return f"bridge_port/{name}"

    def CreateNetworkFunction(self, ingress_mac, egress_mac, ingress_pr_mac, egress_pr_mac):
        macs = [self._seen(x) for x in (ingress_mac, egress_mac, ingress_pr_mac, egress_pr_mac)]
        self.rules += add_nf_p4_rules(self.vf_macs, *macs)

    def AddPeerToPeer(self):
        self.rules += add_peer_to_peer_p4_rules(self.vf_macs)
