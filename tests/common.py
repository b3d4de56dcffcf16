"""Shared test plumbing: the oracle bindings, and scenario builders that stage the SAME rule
set into the CUDA library (through its C ABI), the C++ oracle and the Python oracle."""
from __future__ import annotations

import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

sfc = importlib.import_module("dpu-operator_b200")
consts = sfc.consts

sys.path.insert(0, os.path.join(ROOT, "oracle"))
import orc as _orc  # noqa: E402  (oracle bindings: no product code behind them)
from orc import ORACLE_SO, Oracle, OracleGen, olib  # noqa: E402,F401
from py_oracle import PyOracle  # noqa: E402

assert _orc.FLOW_KEY_DTYPE == sfc.FLOW_KEY_DTYPE and _orc.FLOW_ACTION_DTYPE == sfc.FLOW_ACTION_DTYPE
assert _orc.ACL_RULE_DTYPE == sfc.ACL_RULE_DTYPE and _orc.GEN_SPEC_DTYPE == sfc.GEN_SPEC_DTYPE
assert _orc.CNT_TOTAL == consts.CNT_TOTAL


def _prio(in_port, out_port, mac):
    if mac is None:
        return consts.PRIO_WILDCARD
    return consts.PRIO_HAIRPIN if in_port == out_port else consts.PRIO_MAC


class Scenario:
    """Stages every rule into all sinks at once.  `dp` is a Dataplane (online or offline) and
    is the authority for ids; the oracles receive the ids it returns."""

    def __init__(self, dp, with_py=False):
        self.dp = dp
        self.orc = Oracle()
        self.py = PyOracle() if with_py else None
        self.l2_rules = []  # (in, out, mac, id) in add order, for replay checks

    # ---- rules ----
    def l2_add(self, in_port, out_port, mac=None):
        rid = self.dp.l2_rule_add(in_port, out_port, mac)
        self.orc.l2_add(in_port, out_port, mac, _prio(in_port, out_port, mac), rid)
        if self.py:
            self.py.l2_add(in_port, out_port, None if mac is None else bytes(mac), _prio(in_port, out_port, mac), rid)
        self.l2_rules.append((in_port, out_port, mac, rid))
        return rid

    def l2_del(self, in_port, mac=None):
        self.dp.l2_rule_del(in_port, mac)
        self.orc.l2_del(in_port, mac)
        if self.py:
            self.py.l2_del(in_port, None if mac is None else bytes(mac))

    def chain_set(self, ins, outs):
        self.dp.chain_set(ins, outs)
        self.orc.chain_set(ins, outs)
        if self.py:
            self.py.chain_set(list(ins), list(outs))

    def flows_add(self, keys, acts):
        first = self.dp.flow_add_bulk(keys, acts)
        self.orc.flow_add(keys, acts, first)
        if self.py:
            for i in range(len(keys)):
                k, a = keys[i], acts[i]
                w = 4 if k["ip_ver"] == 4 else 16
                self.py.flow_add(int(k["ip_ver"]), int(k["vlan"]), int(k["proto"]),
                                 int.from_bytes(bytes(k["src"][:w]), "big"), int.from_bytes(bytes(k["dst"][:w]), "big"),
                                 int(k["sport"]), int(k["dport"]), int(a["action"]), int(a["out_port"]), first + i)
        return first

    def lpm_add(self, af, prefix, length, port):
        pid = self.dp.lpm_add(af, prefix, length, port)
        self.orc.lpm_add(af, prefix, length, port, pid)
        if self.py:
            self.py.lpm_add(af, bytes(prefix).ljust(16, b"\0"), length, port, pid)
        return pid

    def acl_add(self, rule):
        rid = self.dp.acl_add(rule)
        self.orc.acl_add(rule, rid)
        if self.py:
            r = np.ascontiguousarray(rule, dtype=sfc.ACL_RULE_DTYPE).reshape(1)[0]
            d = {n: (bytes(r[n]) if r[n].ndim else int(r[n])) for n in sfc.ACL_RULE_DTYPE.names}
            self.py.acl_add(d, rid)
        return rid

    # ---- the SFC steering set, generalised to a chain of NFs ------------------------------
    def build_chain(self, n_vfs=8, n_nfs=1):
        """Marvell VSP rule set (marvell/main.go:382-403, 519-560) for n_vfs host VFs and a chain of
        n_nfs network functions; n_nfs == 1 is exactly the reference's 3*|VF|+2 rules."""
        dp = self.dp
        wire = dp.port_add("rpm0")
        vfs = [dp.port_add(f"vf0-{i}") for i in range(n_vfs)]
        nf_in = [dp.port_add(f"nf{i}-in") for i in range(n_nfs)]
        nf_out = [dp.port_add(f"nf{i}-out") for i in range(n_nfs)]
        self.chain_set(nf_in, nf_out)
        for v, port in enumerate(vfs):
            mac = sfc.gen_mac(v)
            self.l2_add(port, nf_in[0], None)
            self.l2_add(nf_in[0], port, mac)
            self.l2_add(nf_out[-1], nf_out[-1], mac)
            for i in range(1, n_nfs):  # towards the VF: NF[i].in -> NF[i-1].out by dl_dst
                self.l2_add(nf_in[i], nf_out[i - 1], mac)
        for i in range(n_nfs - 1):  # away from the VF
            self.l2_add(nf_out[i], nf_in[i + 1], None)
        self.l2_add(nf_out[-1], wire, None)
        self.l2_add(wire, nf_out[-1], None)
        return dict(wire=wire, vfs=vfs, nf_in=nf_in, nf_out=nf_out, n_ports=1 + n_vfs + 2 * n_nfs)


# ---------------------------------------------------------------- BASELINE.json configs --------
def config2_spec(**kw):
    """1K exact 5-tuple rules, IPv4, 90% hits."""
    d = dict(n_flows=1024, n_ports=8, n_out_ports=8, hit_permille=900, nonip_permille=0, v6_permille=0,
             vlan_permille=0, zipf=0, n_macs=8)
    d.update(kw)
    return sfc.make_gen_spec(**d)


def config3_spec(n_flows=1 << 20, n_ports=17, **kw):
    """1M-flow table, Zipf traffic, 5% misses, 10% non-IP, 4-NF chain on 8 VFs (config 3)."""
    d = dict(n_flows=n_flows, n_ports=n_ports, n_out_ports=n_ports, hit_permille=947, nonip_permille=100,
             v6_permille=200, vlan_permille=250, zipf=1, n_macs=12)
    d.update(kw)
    return sfc.make_gen_spec(**d)


def config4_spec(n_pfx4=80000, n_pfx6=20000, n_acl=1024, **kw):
    """LPM + ACL, 70/25/5 v4/v6/non-IP, no flow table; traffic enters on ports without L2 rules."""
    d = dict(n_flows=0, n_ports=4, port_base=32, n_out_ports=16, hit_permille=0, nonip_permille=50,
             v6_permille=263, vlan_permille=100, n_macs=8, n_pfx4=n_pfx4, n_pfx6=n_pfx6, lpm_permille=900,
             n_acl=n_acl)
    d.update(kw)
    return sfc.make_gen_spec(**d)


def install_flows(sc: Scenario, spec, chunk=1 << 18):
    n = int(spec["n_flows"][0])
    for first in range(0, n, chunk):
        keys, acts = sfc.gen_flows(spec, first, min(chunk, n - first))
        got = sc.flows_add(keys, acts)
        assert got == first


def install_lpm_acl(sc: Scenario, spec):
    for af, n in ((4, int(spec["n_pfx4"][0])), (6, int(spec["n_pfx6"][0]))):
        for p, ln, port in sfc.gen_prefixes(spec, af, n):
            sc.lpm_add(af, p, ln, port)
    rules = sfc.gen_acl_rules(spec, int(spec["n_acl"][0]))
    for r in rules:
        sc.acl_add(r)
    return rules
