"""Second rule dialect: the P4Runtime rule strings the Intel IPU vendor plugin
emits, loaded unchanged.

The Intel VSP (vendored ``ipu-plugin/pkg/p4rtclient/p4rtclient.go``) programs its FXP pipeline by
exec'ing ``p4rt-ctl add-entry br0 <table> "<match>,action=<action>(<params>)"`` once per rule.  This
module restates the part of ``cmd/intelvsp/p4rt-ctl`` that turns such a string into a P4Runtime
``TableEntry`` —

  * the grammar:   ``parse_match_key`` / ``parse_match_key_ternary`` / ``parse_action`` (p4rt-ctl:887-970),
                   dispatch on ",priority=" as in ``p4ctl_add_entry`` (:1083-1100);
  * the encoding:  ``encode`` / ``encodeNum`` / ``encodeMac`` / ``encodeIPv4`` / ``encodeIPv6`` (:200-293)
                   — big-endian, ceil(bitwidth / 8) bytes;
  * the typing:    EXACT / LPM / TERNARY per p4info match field (:380-404), priority kept only if some
                   field is ternary (:455-534);

— against the table and action schemas of the reference's p4info
(``cmd/intelvsp/fxp-net_linux-networking/fxp-net_linux-networking.p4info.txt``, shipped here as
``data/fxp_p4info_schema.json``).  tests/test_p4rt.py checks every entry, down to the serialized
protobuf bytes, against fixtures produced by running the reference's own p4rt-ctl code.

``Importer`` then maps the tables whose meaning is expressible in the B200 pipeline onto the C ABI:

  ipv4_table / ipv6_table (LPM on the destination -> next-hop id) + nexthop_table (id -> egress port)
        -> ``sfc_lpm_add``;
  source_port_to_bridge_map (port -> bridge) + l2_fwd_rx_table / l2_fwd_tx_table ((bridge, dst_mac) -> port)
        -> ``sfc_l2_rule_add(in_port, port, mac)`` for every member port of the bridge;
  source_port_to_pr_map (port -> port)      -> ``sfc_l2_rule_add(in_port, port, NULL)``.

Everything else (VSI metadata plumbing, LAG, mirroring, IPsec) is parsed, validated, kept and
reported as *unmapped*: the P4 source of the pipeline is not in the reference, so its apply order —
and with it the meaning of those tables for a packet — is not recoverable.
"""
from __future__ import annotations

import ipaddress
import json
import os
import re
import socket
from dataclasses import dataclass, field

_SCHEMA_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "fxp_p4info_schema.json")
_MAC = re.compile(r"^([\da-fA-F]{2}:){5}([\da-fA-F]{2})$")     # p4rt-ctl:195
_IPV4 = re.compile(r"^(\d{1,3}\.){3}(\d{1,3})$")               # p4rt-ctl:219


class P4Error(ValueError):
    pass


def load_schema(path: str = _SCHEMA_PATH) -> dict:
    return json.load(open(path))


# ------------------------------------------------------------------ encoding (p4rt-ctl:200-293)
def _is_ipv6(s: str) -> bool:
    try:
        ipaddress.IPv6Address(s)
        return True
    except ValueError:
        return False


def encode_num(number: int, bitwidth: int) -> bytes:
    if number >= 2 ** bitwidth:
        raise P4Error(f"Number, {number}, does not fit in {bitwidth} bits")
    return number.to_bytes((bitwidth + 7) // 8, "big")


def encode(x, bitwidth: int) -> bytes:
    """Infer the type of ``x`` and encode it the way p4rt-ctl's ``encode`` does."""
    byte_len = (bitwidth + 7) // 8
    if isinstance(x, (list, tuple)) and len(x) == 1:
        x = x[0]
    if isinstance(x, str):
        if _MAC.match(x):
            out = bytes.fromhex(x.replace(":", ""))
        elif _IPV4.match(x):
            out = socket.inet_aton(x)
        elif _is_ipv6(x):
            out = socket.inet_pton(socket.AF_INET6, x)
        elif x.isdigit():
            out = encode_num(int(x), bitwidth)
        elif x[:2] == "0x" and all(c in "0123456789abcdefABCDEF" for c in x[2:]):
            out = encode_num(int(x[2:], 16), bitwidth)
        else:
            raise P4Error(f"cannot encode {x!r}")
    elif isinstance(x, int):
        out = encode_num(x, bitwidth)
    else:
        raise P4Error(f"Encoding objects of {type(x)!r} is not supported")
    if len(out) != byte_len:
        raise P4Error(f"{x!r} encodes to {len(out)} bytes, field is {byte_len}")
    return out


# ------------------------------------------------------------------ grammar (p4rt-ctl:887-970)
def parse_match_key(key: str) -> dict:
    out = {}
    for mk in key.split(","):
        name, _, val = mk.partition("=")
        if "/" in val:
            v, _, l = val.partition("/")
            out[name] = (v, int(l) if l.isdigit() else l)  # "/24" prefix length, "/0x7ff" ternary mask
        else:
            out[name] = val
    return out


def parse_action(action: str, schema: dict):
    name, _, rest = action.partition("(")
    if not rest:
        return name, None
    data = rest.split(")")[0].split(",")
    if name not in schema["actions"]:
        raise P4Error(f"unknown action {name!r}")
    params = schema["actions"][name]["params"]
    if data[0].find("=") != -1:  # name=value form: unspecified parameters default to 0 (p4rt-ctl:944-955)
        values = ["0"] * len(params)
        for p in params:
            for item in data:
                k, _, v = item.partition("=")
                if k == p["name"]:
                    values[p["id"] - 1] = v
        data = values

    def to_int(a: str) -> int:
        if _IPV4.match(a):
            return int.from_bytes(socket.inet_aton(a), "big")
        if _is_ipv6(a):
            return int.from_bytes(socket.inet_pton(socket.AF_INET6, a), "big")
        if _MAC.match(a):
            return int(a.replace(":", ""), 16)
        return int(a, 0) if a.find("0x") != -1 else int(a)

    return name, [to_int(a) for a in data]


@dataclass
class Match:
    field_id: int
    name: str
    kind: str  # exact | lpm | ternary
    value: bytes
    mask: bytes | None = None
    prefix_len: int | None = None


@dataclass
class TableEntry:
    table: str
    table_id: int
    match: list[Match]
    priority: int
    action: str
    action_id: int
    params: list[tuple[int, str, bytes]] = field(default_factory=list)  # (param id, name, value)

    def param(self, name: str) -> int:
        for _, n, v in self.params:
            if n == name:
                return int.from_bytes(v, "big")
        raise KeyError(name)

    def field(self, name: str) -> Match:
        for m in self.match:
            if m.name == name:
                return m
        raise KeyError(name)

    # ---- protobuf wire format of p4.v1.TableEntry, hand-encoded (no generated code needed) ----
    def serialize(self) -> bytes:
        def varint(n):
            out = bytearray()
            while True:
                b = n & 0x7F
                n >>= 7
                out.append(b | (0x80 if n else 0))
                if not n:
                    return bytes(out)

        def ld(fno, payload):
            return varint((fno << 3) | 2) + varint(len(payload)) + payload

        def vi(fno, n):
            return varint(fno << 3) + varint(n) if n else b""

        out = vi(1, self.table_id)
        for m in self.match:
            body = vi(1, m.field_id)
            if m.kind == "exact":
                body += ld(2, ld(1, m.value))
            elif m.kind == "ternary":
                body += ld(3, ld(1, m.value) + ld(2, m.mask))
            elif m.kind == "lpm":
                body += ld(4, ld(1, m.value) + vi(2, m.prefix_len))
            out += ld(2, body)
        act = vi(1, self.action_id) + b"".join(ld(4, vi(2, pid) + ld(3, val)) for pid, _, val in self.params)
        out += ld(3, ld(1, act))
        out += vi(4, self.priority)
        return out


def parse_entry(table: str, flow: str, schema: dict | None = None) -> TableEntry:
    """``p4rt-ctl add-entry br0 <table> <flow>`` up to and including ``buildTableEntry``."""
    schema = schema or load_schema()
    if table not in schema["tables"]:
        raise P4Error(f"unknown table {table!r}")
    t = schema["tables"][table]
    if ",action=" not in flow:
        raise P4Error("flow has no ',action='")
    mk, act = flow.split(",action=")[0], flow.split(",action=")[1]
    priority = None
    if flow.find("priority") != -1:
        if ",priority=" not in mk:
            raise P4Error("priority must follow the match key")
        mk, priority = mk.split(",priority=")
    keys = parse_match_key(mk)
    action, data = parse_action(act, schema)
    if action not in schema["actions"]:
        raise P4Error(f"unknown action {action!r}")
    if action not in t["actions"]:
        raise P4Error(f"action {action!r} is not valid for table {table!r}")
    fields = {f["name"]: f for f in t["match_fields"]}
    matches, any_ternary = [], False
    for name, value in keys.items():
        if name not in fields:
            raise P4Error(f"table {table!r} has no match field {name!r}")
        f = fields[name]
        if f["kind"] == "exact":
            matches.append(Match(f["id"], name, "exact", encode(value, f["bitwidth"])))
        elif f["kind"] == "lpm":
            if not isinstance(value, tuple):
                raise P4Error(f"LPM field {name!r} needs value/prefix_len")
            matches.append(Match(f["id"], name, "lpm", encode(value[0], f["bitwidth"]), prefix_len=int(value[1])))
        elif f["kind"] == "ternary":
            any_ternary = True
            if priority is None:
                raise P4Error("For ternary match_type, priority field is needed")  # p4rt-ctl:486-489
            if not isinstance(value, tuple):
                raise P4Error(f"ternary field {name!r} needs value/mask")
            matches.append(Match(f["id"], name, "ternary", encode(value[0], f["bitwidth"]),
                                 mask=encode(str(value[1]), f["bitwidth"])))
        else:
            raise P4Error(f"Unsupported match type {f['kind']!r}")
    a = schema["actions"][action]
    params = []
    if data is not None:
        if len(data) != len(a["params"]):
            raise P4Error(f"action {action!r} takes {len(a['params'])} parameters, got {len(data)}")
        params = [(p["id"], p["name"], encode(int(data[i]), p["bitwidth"])) for i, p in enumerate(a["params"])]
    return TableEntry(table, t["id"], matches, int(priority) if (priority is not None and any_ternary) else 0, action, a["id"], params)


# ------------------------------------------------------------------ mapping onto the B200 dataplane
_C = "linux_networking_control."


class Importer:
    """Collects entries, then ``apply(dp)`` stages the mappable ones through the C ABI (``dp`` is a
    ``Dataplane``; port numbers of the Intel pipeline are used as port ids one to one)."""

    def __init__(self, schema: dict | None = None):
        self.schema = schema or load_schema()
        self.entries: list[TableEntry] = []
        self.unmapped: list[TableEntry] = []

    def add(self, table: str, flow: str) -> TableEntry:
        e = parse_entry(table, flow, self.schema)
        self.entries.append(e)
        return e

    def apply(self, dp, vsi_mac: dict | None = None) -> dict:
        """Stage what the entries mean through the C ABI.  Mapping (in_port = the VSI a frame comes from, output =
        the vport the action names; the compiled pipeline's apply order is not in the repository, so this is the
        reading the Intel plugin's own rule sets imply):
          ipv4_table / ipv6_table + nexthop_table      -> LPM prefix -> the next hop's egress port
          source_port_to_bridge_map + l2_fwd_rx/tx     -> (in_port of the bridge, dst MAC) -> port
          tx_source_port (vsi -> source port)          -> names the VSI behind a source port (no rule by itself)
          source_port_to_pr_map (source port -> port)  -> in_port wildcard rule (priority 10 shape)
          tx_acc_vsi (vsi -> l2_fwd_and_bypass_bridge) -> in_port wildcard rule
          vsi_to_vsi_loopback (vsi, target_vsi -> port)-> (in_port, dst MAC of target_vsi) rule; `vsi_mac` supplies the
                                                          MAC of a VSI (the plugin knows it, the rule string does not)
        Everything else (LAG, mirror, IPsec, tunnel tables) is kept and reported as unmapped."""
        vsi_mac = vsi_mac or {}
        nexthop_port, routes, bridge_of, l2, wild, vsi_of_sp, loops = {}, [], {}, [], [], {}, []
        self.unmapped = []
        for e in self.entries:
            short = e.table[len(_C):] if e.table.startswith(_C) else e.table
            if short == "nexthop_table" and e.action.endswith("set_nexthop_info_dmac"):
                nexthop_port[int.from_bytes(e.field("user_meta.cmeta.nexthop_id").value, "big")] = e.param("egress_port")
            elif short in ("ipv4_table", "ipv6_table") and e.action.endswith("_set_nexthop_id"):
                m = e.field("ipv4_dst_match" if short == "ipv4_table" else "ipv6_dst_match")
                routes.append((4 if short == "ipv4_table" else 6, m.value, m.prefix_len, e.param("nexthop_id")))
            elif short == "source_port_to_bridge_map" and e.action.endswith("set_bridge_id"):
                sp = e.field("user_meta.cmeta.source_port")
                if int.from_bytes(sp.mask, "big") == 0xFFFF:  # one port per entry, as the VSP writes them
                    bridge_of[int.from_bytes(sp.value, "big")] = e.param("bridge_id")
                else:
                    self.unmapped.append(e)
            elif short in ("l2_fwd_rx_table", "l2_fwd_tx_table") and e.action.endswith(".l2_fwd"):
                l2.append((int.from_bytes(e.field("user_meta.pmeta.bridge_id").value, "big"), e.field("dst_mac").value, e.param("port")))
            elif short == "source_port_to_pr_map" and e.action.endswith("fwd_to_vsi"):
                wild.append(("sp", int.from_bytes(e.field("user_meta.cmeta.source_port").value, "big"), e.param("port")))
            elif short == "tx_source_port" and e.action.endswith("set_source_port"):
                v = e.field("vmeta.common.vsi")
                if int.from_bytes(v.mask, "big") == 0x7FF:  # one VSI per entry, as the VSP writes them
                    vsi_of_sp[e.param("source_port")] = int.from_bytes(v.value, "big")
                else:
                    self.unmapped.append(e)
            elif short == "tx_acc_vsi" and e.action.endswith("l2_fwd_and_bypass_bridge"):
                wild.append(("vsi", int.from_bytes(e.field("vmeta.common.vsi").value, "big"), e.param("port")))
            elif short == "vsi_to_vsi_loopback" and e.action.endswith("fwd_to_vsi"):
                target = int.from_bytes(e.field("target_vsi").value, "big")
                if target in vsi_mac:
                    loops.append((int.from_bytes(e.field("vmeta.common.vsi").value, "big"), bytes(vsi_mac[target]), e.param("port")))
                else:
                    self.unmapped.append(e)  # nobody told us which MAC this VSI answers to
            else:
                self.unmapped.append(e)
        done = dict(lpm=0, l2_exact=0, l2_wildcard=0, unresolved_routes=0)
        for af, value, plen, nh in routes:
            if nh not in nexthop_port:
                done["unresolved_routes"] += 1  # points at a next hop nobody programmed (or that drops)
                continue
            dp.lpm_add(af, value, plen, nexthop_port[nh])
            done["lpm"] += 1
        for bridge, mac, port in l2:
            for in_port, b in sorted(bridge_of.items()):
                if b == bridge and in_port != port:
                    dp.l2_rule_add(in_port, port, mac)
                    done["l2_exact"] += 1
        for in_port, mac, port in loops:
            dp.l2_rule_add(in_port, port, mac)
            done["l2_exact"] += 1
        for kind, key, port in wild:
            in_port = vsi_of_sp.get(key, key) if kind == "sp" else key  # a source port stands for the VSI that tx_source_port gave it
            dp.l2_rule_add(in_port, port, None)
            done["l2_wildcard"] += 1
        done["unmapped"] = len(self.unmapped)
        done["unmapped_tables"] = sorted({e.table[len(_C):] if e.table.startswith(_C) else e.table for e in self.unmapped})
        return done
