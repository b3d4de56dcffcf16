"""ctypes view of oracle/libsfc_oracle.so: the CPU oracle (sfc_oracle.cpp) and its own writing of the synthetic
workload (workload.cpp).  TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  It imports nothing from the product package and never loads libb200sfc.so:
the struct layouts below restate include/b200sfc.h (tests/test_oracle.py holds them to the product's)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libsfc_oracle.so")

FLOW_KEY_DTYPE = np.dtype([("src", "u1", 16), ("dst", "u1", 16), ("sport", "<u2"), ("dport", "<u2"), ("vlan", "<u2"),
                           ("proto", "u1"), ("ip_ver", "u1")])
FLOW_ACTION_DTYPE = np.dtype([("action", "u1"), ("reserved", "u1"), ("out_port", "<u2")])
ACL_RULE_DTYPE = np.dtype([("src", "u1", 16), ("src_mask", "u1", 16), ("dst", "u1", 16), ("dst_mask", "u1", 16),
                           ("in_port", "<u2"), ("in_port_mask", "<u2"), ("vlan", "<u2"), ("vlan_mask", "<u2"),
                           ("sport", "<u2"), ("sport_mask", "<u2"), ("dport", "<u2"), ("dport_mask", "<u2"),
                           ("proto", "u1"), ("proto_mask", "u1"), ("ip_ver", "u1"), ("verdict", "u1"), ("priority", "<u4")])
GEN_SPEC_DTYPE = np.dtype([("seed_pkts", "<u8"), ("seed_rules", "<u8"), ("seed_lpm", "<u8"), ("seed_acl", "<u8"),
                           ("n_flows", "<u4"), ("n_ports", "<u4"), ("port_base", "<u4"), ("n_out_ports", "<u4"),
                           ("hit_permille", "<u4"), ("nonip_permille", "<u4"), ("v6_permille", "<u4"), ("vlan_permille", "<u4"),
                           ("zipf", "<u4"), ("n_macs", "<u4"), ("n_pfx4", "<u4"), ("n_pfx6", "<u4"), ("lpm_permille", "<u4"),
                           ("n_acl", "<u4")])
CNT_TOTAL = 16 + 4096 + 16384  # SFC_CNT_TOTAL
SEEDS = dict(seed_rules=0x5FC00001, seed_pkts=0x5FC00002, seed_lpm=0x5FC00003, seed_acl=0x5FC00004)


def make_gen_spec(**kw) -> np.ndarray:
    s = np.zeros(1, dtype=GEN_SPEC_DTYPE)
    for k, v in SEEDS.items():
        s[k] = v
    s["n_ports"], s["n_out_ports"], s["n_macs"] = 8, 8, 8
    for k, v in kw.items():
        s[k] = v
    return s




def _load_oracle():
    if not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-C", HERE])
    L = C.CDLL(ORACLE_SO)
    vp = C.c_void_p
    L.orc_create.restype = vp
    L.orc_destroy.argtypes = [vp]
    L.orc_clear.argtypes = [vp]
    L.orc_l2_add.argtypes = [vp, C.c_uint16, C.c_uint16, vp, C.c_uint32, C.c_uint32]
    L.orc_l2_del.argtypes = [vp, C.c_uint16, vp]
    L.orc_chain_set.argtypes = [vp, vp, vp, C.c_size_t]
    L.orc_flow_add.argtypes = [vp, vp, vp, C.c_size_t, C.c_uint32]
    L.orc_flow_del.argtypes = [vp, vp]
    L.orc_lpm_add.argtypes = [vp, C.c_int, vp, C.c_int, C.c_uint16, C.c_uint32]
    L.orc_acl_add.argtypes = [vp, vp, C.c_uint32]
    L.orc_classify.argtypes = [vp, vp, vp, vp, C.c_size_t, C.c_int, vp]
    L.orc_parse.argtypes = [vp, vp, vp, vp, vp]
    # oracle/workload.cpp: the synthetic workload written a second time
    L.orc_gen_packets.argtypes = [vp, C.c_uint64, C.c_size_t, vp, vp]
    L.orc_gen_flows.argtypes = [vp, C.c_uint32, C.c_size_t, vp, vp]
    L.orc_gen_mac.argtypes = [C.c_uint32, vp]
    L.orc_gen_prefix.argtypes = [vp, C.c_int, C.c_uint32, vp, C.POINTER(C.c_int), C.POINTER(C.c_uint16)]
    L.orc_gen_acl_rule.argtypes = [vp, C.c_uint32, vp]
    return L


olib = _load_oracle()


def _macbuf(mac):
    return None if mac is None else C.cast(C.create_string_buffer(bytes(mac), 6), C.c_void_p)


class Oracle:
    """ctypes view of oracle/libsfc_oracle.so."""

    def __init__(self):
        self._o = C.c_void_p(olib.orc_create())

    def __del__(self):
        if getattr(self, "_o", None):
            olib.orc_destroy(self._o)
            self._o = None

    def l2_add(self, in_port, out_port, mac, priority, rid):
        olib.orc_l2_add(self._o, in_port, out_port, _macbuf(mac), priority, rid)

    def l2_del(self, in_port, mac=None):
        olib.orc_l2_del(self._o, in_port, _macbuf(mac))

    def chain_set(self, ins, outs):
        a = np.ascontiguousarray(ins, dtype=np.uint16)
        b = np.ascontiguousarray(outs, dtype=np.uint16)
        olib.orc_chain_set(self._o, a.ctypes.data, b.ctypes.data, a.size)

    def flow_add(self, keys, acts, first_id):
        keys = np.ascontiguousarray(keys, dtype=FLOW_KEY_DTYPE)
        acts = np.ascontiguousarray(acts, dtype=FLOW_ACTION_DTYPE)
        olib.orc_flow_add(self._o, keys.ctypes.data, acts.ctypes.data, keys.size, first_id)

    def flow_del(self, key):
        k = np.ascontiguousarray(key, dtype=FLOW_KEY_DTYPE).reshape(1)
        olib.orc_flow_del(self._o, k.ctypes.data)

    def lpm_add(self, af, prefix, length, port, pid):
        buf = C.create_string_buffer(bytes(prefix).ljust(16, b"\0"), 16)
        olib.orc_lpm_add(self._o, af, C.cast(buf, C.c_void_p), length, port, pid)

    def acl_add(self, rule, rid):
        r = np.ascontiguousarray(rule, dtype=ACL_RULE_DTYPE).reshape(1)
        olib.orc_acl_add(self._o, r.ctypes.data, rid)

    def classify(self, hdr, in_port, threads=1, counters=False):
        hdr = np.ascontiguousarray(hdr, dtype=np.uint8)
        in_port = np.ascontiguousarray(in_port, dtype=np.uint16)
        out = np.empty(in_port.size, dtype=np.uint64)
        cnt = np.zeros(CNT_TOTAL, dtype=np.uint64) if counters else None
        olib.orc_classify(self._o, hdr.ctypes.data, in_port.ctypes.data, out.ctypes.data, in_port.size, threads,
                          cnt.ctypes.data if counters else None)
        return (out, cnt) if counters else out

    @staticmethod
    def parse(hdr_row):
        h = np.ascontiguousarray(hdr_row, dtype=np.uint8)
        f = np.zeros(5, dtype=np.uint32)
        s, d, m = np.zeros(16, np.uint8), np.zeros(16, np.uint8), np.zeros(6, np.uint8)
        olib.orc_parse(h.ctypes.data, f.ctypes.data, s.ctypes.data, d.ctypes.data, m.ctypes.data)
        return dict(vlan=int(f[0]), ver=int(f[1]), proto=int(f[2]), sport=int(f[3]), dport=int(f[4]),
                    src=bytes(s), dst=bytes(d), dmac=bytes(m))


class OracleGen:
    """The oracle's own writing of the synthetic workload (oracle/workload.cpp); same call shapes as the
    product's sfc.gen_* helpers, no product code behind it."""

    @staticmethod
    def packets(spec, first, n):
        hdr = np.empty((n, 64), dtype=np.uint8)
        ports = np.empty(n, dtype=np.uint16)
        olib.orc_gen_packets(spec.ctypes.data, first, n, hdr.ctypes.data, ports.ctypes.data)
        return hdr, ports

    @staticmethod
    def flows(spec, first, n):
        keys = np.zeros(n, dtype=FLOW_KEY_DTYPE)
        acts = np.zeros(n, dtype=FLOW_ACTION_DTYPE)
        olib.orc_gen_flows(spec.ctypes.data, first, n, keys.ctypes.data, acts.ctypes.data)
        return keys, acts

    @staticmethod
    def mac(i):
        buf = C.create_string_buffer(6)
        olib.orc_gen_mac(i, C.cast(buf, C.c_void_p))
        return buf.raw

    @staticmethod
    def prefixes(spec, af, n):
        out = []
        buf, ln, port = C.create_string_buffer(16), C.c_int(), C.c_uint16()
        for j in range(n):
            olib.orc_gen_prefix(spec.ctypes.data, af, j, C.cast(buf, C.c_void_p), C.byref(ln), C.byref(port))
            out.append((buf.raw[: 4 if af == 4 else 16], ln.value, port.value))
        return out

    @staticmethod
    def acl_rules(spec, n):
        rules = np.zeros(n, dtype=ACL_RULE_DTYPE)
        for j in range(n):
            olib.orc_gen_acl_rule(spec.ctypes.data, j, rules[j:].ctypes.data)
        return rules


