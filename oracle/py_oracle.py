"""py_oracle.py — a second, independent restatement of the classification semantics, in plain
Python (linear scans, Python ints, no shared code with sfc_oracle.cpp or the CUDA library).

TEST INFRASTRUCTURE ONLY; PARITY UNPINNED (see the header of sfc_oracle.cpp: the reference has
no per-packet golden vectors).  Its job is to catch a mistake that the C++ oracle and the CUDA
kernel could share because the same person wrote both from the same notes: tests run the two
oracles against each other on small seeded cases and on the hand-built fixtures in tests/golden.

Reference lines restated: rule shapes and priorities of
internal/daemon/vendor-specific-plugins/marvell/ovs-dp/ovsdp.go:113-142; ternary/priority
semantics of cmd/intelvsp/p4rt-ctl:380-404,455-534; LPM tables of
cmd/intelvsp/fxp-net_linux-networking/fxp-net_linux-networking.p4info.txt:807,868; wire byte
order as in dpu-cni/pkgs/sriovutils/packet.go:41-66.
"""
from __future__ import annotations

ACT_TO_CP, ACT_FORWARD, ACT_DROP = 0, 1, 2
NF_NONE = 0xFF
ST_MISS, ST_ACL, ST_FLOW, ST_L2X, ST_L2W, ST_LPM4, ST_LPM6 = range(7)


def result(action, nf, port, stage, index):
    return (action << 56) | (nf << 48) | (port << 32) | (stage << 28) | index


def parse(h: bytes):
    """-> dict(dmac, vlan, ver, proto, src, dst, sport, dport); src/dst are ints."""
    assert len(h) == 64
    k = dict(dmac=bytes(h[0:6]), vlan=0, ver=0, proto=0, src=0, dst=0, sport=0, dport=0)
    et = int.from_bytes(h[12:14], "big")
    off = 14
    if et == 0x8100:
        k["vlan"] = int.from_bytes(h[14:16], "big") & 0xFFF
        et = int.from_bytes(h[16:18], "big")
        off = 18
    if et == 0x0800:
        ver, ihl = h[off] >> 4, h[off] & 15
        if ver != 4 or ihl < 5:
            return k
        k["ver"], k["proto"] = 4, h[off + 9]
        k["src"] = int.from_bytes(h[off + 12 : off + 16], "big")
        k["dst"] = int.from_bytes(h[off + 16 : off + 20], "big")
        frag = int.from_bytes(h[off + 6 : off + 8], "big") & 0x1FFF
        l4 = off + 4 * ihl
        if k["proto"] in (6, 17, 132) and frag == 0 and l4 + 4 <= 64:
            k["sport"] = int.from_bytes(h[l4 : l4 + 2], "big")
            k["dport"] = int.from_bytes(h[l4 + 2 : l4 + 4], "big")
    elif et == 0x86DD:
        if h[off] >> 4 != 6:
            return k
        k["ver"], k["proto"] = 6, h[off + 6]
        k["src"] = int.from_bytes(h[off + 8 : off + 24], "big")
        k["dst"] = int.from_bytes(h[off + 24 : off + 40], "big")
        l4 = off + 40
        if k["proto"] in (6, 17, 132) and l4 + 4 <= 64:
            k["sport"] = int.from_bytes(h[l4 : l4 + 2], "big")
            k["dport"] = int.from_bytes(h[l4 + 2 : l4 + 4], "big")
    return k


class PyOracle:
    def __init__(self):
        self.l2 = []  # (in_port, out_port, mac|None, priority, id)
        self.nfs = []  # (in_port, out_port)
        self.flows = {}  # (ver, vlan, proto, src, dst, sport, dport) -> (action, out_port, id)
        self.lpm = {4: [], 6: []}  # (value_int, len, port, id)
        self.acl = []  # (dict, id)

    # ---- rule staging -------------------------------------------------------------------
    def l2_add(self, in_port, out_port, mac, priority, rid):
        for i, r in enumerate(self.l2):
            if r[0] == in_port and r[2] == mac and r[3] == priority:
                self.l2[i] = (in_port, out_port, mac, priority, r[4])
                return
        self.l2.append((in_port, out_port, mac, priority, rid))

    def l2_del(self, in_port, mac=None):
        self.l2 = [r for r in self.l2 if not (r[0] == in_port and (mac is None or r[2] == mac))]

    def chain_set(self, ins, outs):
        self.nfs = list(zip(ins, outs))

    def flow_add(self, ver, vlan, proto, src, dst, sport, dport, action, out_port, fid):
        self.flows[(ver, vlan, proto, src, dst, sport, dport)] = (action, out_port, fid)

    def lpm_add(self, af, prefix: bytes, length, port, pid):
        bits = 32 if af == 4 else 128
        v = int.from_bytes(prefix[: bits // 8], "big")
        v &= ((1 << length) - 1) << (bits - length) if length else 0
        self.lpm[af].append((v, length, port, pid))

    def acl_add(self, rule: dict, rid):
        self.acl.append((rule, rid))

    # ---- lookup -------------------------------------------------------------------------
    def nf_of(self, port):
        for i, (a, b) in enumerate(self.nfs):
            if port in (a, b):
                return i
        return NF_NONE

    def classify(self, h: bytes, in_port: int) -> int:
        k = parse(h)
        ip = k["ver"] != 0
        if ip:
            best = None
            width = 32 if k["ver"] == 4 else 128
            for r, rid in self.acl:
                if r["ip_ver"] and r["ip_ver"] != k["ver"]:
                    continue
                # addresses of a rule are left-aligned 16-byte strings; IPv4 uses the first 4
                rs = int.from_bytes(bytes(r["src"]), "big") >> (128 - width)
                rsm = int.from_bytes(bytes(r["src_mask"]), "big") >> (128 - width)
                rd = int.from_bytes(bytes(r["dst"]), "big") >> (128 - width)
                rdm = int.from_bytes(bytes(r["dst_mask"]), "big") >> (128 - width)
                if r["ip_ver"] == 0 and (rsm or rdm):
                    raise ValueError("address mask on an any-version rule")
                ok = (
                    (in_port ^ r["in_port"]) & r["in_port_mask"] == 0
                    and (k["vlan"] ^ r["vlan"]) & r["vlan_mask"] == 0
                    and (k["proto"] ^ r["proto"]) & r["proto_mask"] == 0
                    and (k["sport"] ^ r["sport"]) & r["sport_mask"] == 0
                    and (k["dport"] ^ r["dport"]) & r["dport_mask"] == 0
                    and (k["src"] ^ rs) & rsm == 0
                    and (k["dst"] ^ rd) & rdm == 0
                )
                if ok and (best is None or (-r["priority"], rid) < (-best[0]["priority"], best[1])):
                    best = (r, rid)
            if best is not None and best[0]["verdict"] == 1:
                return result(ACT_DROP, NF_NONE, 0, ST_ACL, best[1])
            f = self.flows.get((k["ver"], k["vlan"], k["proto"], k["src"], k["dst"], k["sport"], k["dport"]))
            if f is not None:
                act, port, fid = f
                return result(act, self.nf_of(port) if act == ACT_FORWARD else NF_NONE, port, ST_FLOW, fid)
        cands = [r for r in self.l2 if r[0] == in_port and (r[2] is None or r[2] == k["dmac"])]
        if cands:
            r = min(cands, key=lambda r: (-r[3], r[4]))
            return result(ACT_FORWARD, self.nf_of(r[1]), r[1], ST_L2X if r[2] is not None else ST_L2W, r[4])
        if ip:
            bits = 32 if k["ver"] == 4 else 128
            best = None
            for v, length, port, pid in self.lpm[k["ver"]]:
                if length == 0 or (k["dst"] >> (bits - length)) == (v >> (bits - length)):
                    if best is None or length > best[0]:
                        best = (length, port, pid)
            if best is not None:
                return result(ACT_FORWARD, self.nf_of(best[1]), best[1], ST_LPM4 if k["ver"] == 4 else ST_LPM6, best[2])
        return result(ACT_TO_CP, NF_NONE, 0, ST_MISS, 0)
