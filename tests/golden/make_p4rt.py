"""Regenerate the P4Runtime fixtures from the REFERENCE's own code (run in the build container):

  dpu-operator_b200/data/fxp_p4info_schema.json   names / ids / widths / match kinds of every table and
        action of cmd/intelvsp/fxp-net_linux-networking/fxp-net_linux-networking.p4info.txt
  tests/golden/p4rt_entries.json                  rule strings in the shapes the vendored Intel VSP emits
        (ipu-plugin/pkg/p4rtclient/p4rtclient.go) and the TableEntry each one becomes when pushed
        through the reference's cmd/intelvsp/p4rt-ctl — imported here AS IS (parse_flow,
        parse_flow_ternary, parse_action, P4InfoHelper.buildTableEntry, encode), with `netaddr`
        (absent from this image, used only for valid_ipv6) replaced by the stdlib's ipaddress.

tests/test_p4rt.py holds dpu-operator_b200/p4rt.py to these entries byte for byte.
    python tests/golden/make_p4rt.py
"""
import importlib.machinery
import importlib.util
import ipaddress
import json
import os
import sys
import types
import warnings

REF = "/root/reference/cmd/intelvsp"
sys.path.insert(0, os.path.join(REF, "p4runtime-2023.11.0"))
na = types.ModuleType("netaddr")


def _valid_ipv6(s):
    try:
        ipaddress.IPv6Address(s)
        return True
    except Exception:
        return False


na.valid_ipv6 = _valid_ipv6
sys.modules["netaddr"] = na
warnings.simplefilter("ignore")
loader = importlib.machinery.SourceFileLoader("p4rt_ctl", os.path.join(REF, "p4rt-ctl"))
ctl = importlib.util.module_from_spec(importlib.util.spec_from_loader("p4rt_ctl", loader))
loader.exec_module(ctl)

import google.protobuf.text_format as tf  # noqa: E402
from p4.config.v1 import p4info_pb2  # noqa: E402

info = p4info_pb2.P4Info()
tf.Merge(open(os.path.join(REF, "fxp-net_linux-networking/fxp-net_linux-networking.p4info.txt")).read(), info)
KIND = {p4info_pb2.MatchField.EXACT: "exact", p4info_pb2.MatchField.LPM: "lpm", p4info_pb2.MatchField.TERNARY: "ternary",
        p4info_pb2.MatchField.RANGE: "range", p4info_pb2.MatchField.OPTIONAL: "optional"}
actions = {a.preamble.id: a for a in info.actions}
schema = dict(
    source="cmd/intelvsp/fxp-net_linux-networking/fxp-net_linux-networking.p4info.txt",
    tables={t.preamble.name: dict(id=t.preamble.id, size=t.size,
                                  match_fields=[dict(id=f.id, name=f.name, bitwidth=f.bitwidth, kind=KIND[f.match_type]) for f in t.match_fields],
                                  actions=[actions[r.id].preamble.name for r in t.action_refs]) for t in info.tables},
    actions={a.preamble.name: dict(id=a.preamble.id, params=[dict(id=p.id, name=p.name, bitwidth=p.bitwidth) for p in a.params])
             for a in info.actions},
)
root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
json.dump(schema, open(os.path.join(root, "dpu-operator_b200", "data", "fxp_p4info_schema.json"), "w"), indent=0, sort_keys=True)

C = "linux_networking_control."
RULES = [  # (table, rule string): shapes of p4rtclient.go :247-253, :647-731, :1040-1085 plus the LPM / next-hop tables
    (C + "tx_source_port", f"vmeta.common.vsi=5/0x7ff,priority=1,action={C}set_source_port(21)"),
    (C + "tx_acc_vsi", f"vmeta.common.vsi=9,zero_padding=0,action={C}l2_fwd_and_bypass_bridge(21)"),
    (C + "vsi_to_vsi_loopback", f"vmeta.common.vsi=9,target_vsi=5,action={C}fwd_to_vsi(21)"),
    (C + "vsi_to_vsi_loopback", f"vmeta.common.vsi=5,target_vsi=9,action={C}fwd_to_vsi(25)"),
    (C + "source_port_to_pr_map", f"user_meta.cmeta.source_port=21,zero_padding=0,action={C}fwd_to_vsi(25)"),
    (C + "source_port_to_pr_map", f"user_meta.cmeta.source_port=22,zero_padding=0,action={C}fwd_to_vsi(26)"),
    (C + "source_port_to_bridge_map",
     f"user_meta.cmeta.source_port=0/0xffff,hdrs.vlan_ext[vmeta.common.depth].hdr.vid=0/0xfff,priority=1,action={C}set_bridge_id(bridge_id=77)"),
    (C + "source_port_to_bridge_map",
     f"user_meta.cmeta.source_port=21/0xffff,hdrs.vlan_ext[vmeta.common.depth].hdr.vid=0/0xfff,priority=1,action={C}set_bridge_id(bridge_id=77)"),
    (C + "l2_fwd_rx_table", f"user_meta.pmeta.bridge_id=77,dst_mac=0x000500000314,action={C}l2_fwd(21)"),
    (C + "l2_fwd_rx_table", f"user_meta.pmeta.bridge_id=77,dst_mac=00:11:22:33:44:55,action={C}l2_fwd(25)"),
    (C + "l2_fwd_tx_table", f"user_meta.pmeta.bridge_id=77,dst_mac=02:00:00:00:00:fe,action={C}l2_fwd(0)"),
    (C + "ipv4_table", f"ipv4_table_lpm_root=0,ipv4_dst_match=10.1.2.0/24,action={C}ipv4_set_nexthop_id(7)"),
    (C + "ipv4_table", f"ipv4_table_lpm_root=0,ipv4_dst_match=10.1.0.0/16,action={C}ipv4_set_nexthop_id(8)"),
    (C + "ipv4_table", f"ipv4_table_lpm_root=0,ipv4_dst_match=192.168.7.9/32,action={C}ipv4_set_nexthop_id(0x9)"),
    (C + "ipv6_table", f"ipv6_table_lpm_root=0,ipv6_dst_match=2001:db8:1::/48,action={C}ipv6_set_nexthop_id(7)"),
    (C + "nexthop_table", f"user_meta.cmeta.nexthop_id=7,bit16_zeros=0,action={C}set_nexthop_info_dmac(1,21,0x0011,0x22334455)"),
    (C + "nexthop_table", f"user_meta.cmeta.nexthop_id=8,bit16_zeros=0,action={C}set_nexthop_info_dmac(egress_port=25,router_interface_id=2)"),
    (C + "nexthop_table", f"user_meta.cmeta.nexthop_id=9,bit16_zeros=0,action={C}drop"),
    (C + "ipsec_spd", f"hdrs.ipv4[vmeta.common.depth].dst_ip=10.9.8.7,hdrs.ipv4[vmeta.common.depth].protocol=50,action={C}ipsec_bypass"),
    (C + "tx_lag_table", f"user_meta.cmeta.lag_group_id=0/255,hash=3/7,priority=1,action={C}bypass"),
]
helper = ctl.P4InfoHelper(info)
out = []
for table, flow in RULES:
    prio = None
    if flow.find("priority") != -1:  # the dispatch of p4ctl_add_entry (p4rt-ctl:1083-1100)
        mk, prio, action, data = ctl.parse_flow_ternary(flow, helper)
    else:
        mk, action, data = ctl.parse_flow(flow, helper)
    params = data if (data is None or data == [""]) else {a.name: int(data[i]) for i, a in enumerate(helper.get_action_params(action))}
    te = helper.buildTableEntry(table_name=table, match_fields=mk, action_name=action, action_params=params, priority=prio)
    entry = dict(table=table, rule=flow, table_id=te.table_id, priority=te.priority, action_id=te.action.action.action_id,
                 params=[dict(id=p.param_id, value=p.value.hex()) for p in te.action.action.params], match=[])
    for m in te.match:
        kind = m.WhichOneof("field_match_type")
        d = dict(field_id=m.field_id, kind=kind)
        if kind == "exact":
            d["value"] = m.exact.value.hex()
        elif kind == "lpm":
            d["value"], d["prefix_len"] = m.lpm.value.hex(), m.lpm.prefix_len
        elif kind == "ternary":
            d["value"], d["mask"] = m.ternary.value.hex(), m.ternary.mask.hex()
        entry["match"].append(d)
    entry["wire"] = te.SerializeToString().hex()
    out.append(entry)
json.dump(dict(source="cmd/intelvsp/p4rt-ctl run on the rule shapes of vendored ipu-plugin/pkg/p4rtclient/p4rtclient.go", entries=out),
          open(os.path.join(root, "tests", "golden", "p4rt_entries.json"), "w"), indent=1)
print(f"schema: {len(schema['tables'])} tables, {len(schema['actions'])} actions; {len(out)} golden entries")
