"""Hand-built header records and the rule set they are aimed at.  Nothing here uses the library's
synthetic generator: frames are assembled byte by byte so that parser edge cases are pinned by
construction."""
from __future__ import annotations

import numpy as np

from common import consts, sfc


def eth(dmac: bytes, smac: bytes, ethertype: int, payload: bytes, tags=()) -> bytes:
    f = bytes(dmac) + bytes(smac)
    for tpid, tci in tags:
        f += tpid.to_bytes(2, "big") + tci.to_bytes(2, "big")
    f += ethertype.to_bytes(2, "big") + payload
    return (f + bytes(64))[:64]


def ipv4(src: bytes, dst: bytes, proto: int, l4: bytes, ihl=5, frag=0, version=4) -> bytes:
    h = bytes([(version << 4) | ihl, 0]) + (20 + len(l4)).to_bytes(2, "big") + b"\x12\x34" + frag.to_bytes(2, "big")
    h += bytes([64, proto, 0, 0]) + src + dst
    h += b"\x01" * (4 * (ihl - 5)) if ihl > 5 else b""
    return h + l4


def ipv6(src: bytes, dst: bytes, nh: int, l4: bytes, version=6) -> bytes:
    return bytes([version << 4, 0, 0, 0]) + len(l4).to_bytes(2, "big") + bytes([nh, 64]) + src + dst + l4


def ports(sport, dport) -> bytes:
    return sport.to_bytes(2, "big") + dport.to_bytes(2, "big") + b"\xde\xad\xbe\xef"


def key(ver, src, dst, proto, sport, dport, vlan=0):
    k = np.zeros(1, dtype=sfc.FLOW_KEY_DTYPE)
    k["src"][0, : len(src)] = np.frombuffer(src, np.uint8)
    k["dst"][0, : len(dst)] = np.frombuffer(dst, np.uint8)
    k["sport"], k["dport"], k["vlan"], k["proto"], k["ip_ver"] = sport, dport, vlan, proto, ver
    return k


def act(action, out_port=0):
    a = np.zeros(1, dtype=sfc.FLOW_ACTION_DTYPE)
    a["action"], a["out_port"] = action, out_port
    return a


def acl(ip_ver=0, src=b"", src_len=0, dst=b"", dst_len=0, proto=None, sport=None, dport=None, in_port=None, vlan=None,
        priority=1, deny=True, dport_mask=0xFFFF):
    r = np.zeros(1, dtype=sfc.ACL_RULE_DTYPE)

    def put(field, addr, length):
        if length:
            m = ((1 << length) - 1) << (128 - length)
            mb = m.to_bytes(16, "big")
            ab = (addr + bytes(16))[:16]
            r[field][0] = np.frombuffer(bytes(a & b for a, b in zip(ab, mb)), np.uint8)
            r[field + "_mask"][0] = np.frombuffer(mb, np.uint8)

    put("src", src, src_len)
    put("dst", dst, dst_len)
    r["ip_ver"] = ip_ver
    if proto is not None:
        r["proto"], r["proto_mask"] = proto, 0xFF
    if sport is not None:
        r["sport"], r["sport_mask"] = sport, 0xFFFF
    if dport is not None:
        r["dport"], r["dport_mask"] = dport & dport_mask, dport_mask
    if in_port is not None:
        r["in_port"], r["in_port_mask"] = in_port, 0xFFFF
    if vlan is not None:
        r["vlan"], r["vlan_mask"] = vlan, 0x0FFF
    r["priority"] = priority
    r["verdict"] = consts.ACL_DENY if deny else consts.ACL_PERMIT
    return r


A = bytes([10, 0, 0, 1])
B = bytes([10, 0, 0, 2])
C6 = bytes.fromhex("20010db8000000000000000000000001")
D6 = bytes.fromhex("20010db800010002000300040005beef")
SMAC = bytes.fromhex("02aabbccddee")


def handmade_cases(sc):
    """Stage rules into `sc` (a common.Scenario) and return (hdr[n,64] uint8, in_port[n] uint16)."""
    dp = sc.dp
    topo = sc.build_chain(n_vfs=2, n_nfs=2)  # ports: rpm0=0, vf0-0=1, vf0-1=2, nf0-in=3, nf1-in=4, nf0-out=5, nf1-out=6
    up0 = dp.port_add("uplink0")  # no L2 rules: traffic from here reaches LPM / miss
    up1 = dp.port_add("uplink1")
    vf0, vf1 = topo["vfs"]
    nf_in, nf_out, wire = topo["nf_in"], topo["nf_out"], topo["wire"]
    mac0, mac1 = sfc.gen_mac(0), sfc.gen_mac(1)
    other = bytes.fromhex("0200deadbeef")

    P = []  # (frame, in_port)

    # ---- exact flows --------------------------------------------------------------------
    sc.flows_add(key(4, A, B, 6, 1000, 80), act(consts.ACT_FORWARD, nf_in[1]))          # id 0 -> next_nf 1
    sc.flows_add(key(4, A, B, 17, 1000, 53, vlan=100), act(consts.ACT_FORWARD, wire))   # id 1
    sc.flows_add(key(4, A, B, 6, 0, 0), act(consts.ACT_DROP))                           # id 2: "no ports" key
    sc.flows_add(key(4, A, B, 1, 0, 0), act(consts.ACT_TO_CP))                          # id 3: ICMP
    sc.flows_add(key(4, A, B, 132, 7, 9), act(consts.ACT_FORWARD, vf1))                 # id 4: SCTP
    sc.flows_add(key(6, C6, D6, 6, 443, 50000), act(consts.ACT_FORWARD, nf_in[0]))      # id 5
    sc.flows_add(key(6, C6, D6, 17, 443, 50000, vlan=4095), act(consts.ACT_FORWARD, vf0))  # id 6
    sc.flows_add(key(6, C6, D6, 0, 0, 0), act(consts.ACT_DROP))                         # id 7: hop-by-hop next header
    sc.flows_add(key(4, B, A, 6, 2222, 3333), act(consts.ACT_FORWARD, nf_out[1]))       # id 8: options cases

    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80))), up0))
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 17, ports(1000, 53)), tags=[(0x8100, 0xE064)]), up0))  # PCP/DEI set
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 17, ports(1000, 53)), tags=[(0x8100, 101)]), up0))     # other VID
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 17, ports(1000, 53))), up0))                            # untagged
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80), frag=185)), up0))                   # fragment
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80), frag=0x2000)), up0))                # MF, offset 0
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 1, ports(1000, 80))), up0))                             # ICMP
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 132, ports(7, 9))), up0))                               # SCTP
    for ihl in range(5, 16):  # options push L4 towards / past the end of the 64-byte record
        P.append((eth(other, SMAC, 0x0800, ipv4(B, A, 6, ports(2222, 3333), ihl=ihl)), up0))
        P.append((eth(other, SMAC, 0x0800, ipv4(B, A, 6, ports(2222, 3333), ihl=ihl), tags=[(0x8100, 0)]), up0))
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80), version=5)), up0))   # bad version
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80), ihl=4)), up0))       # bad IHL
    P.append((eth(other, SMAC, 0x86DD, ipv6(C6, D6, 6, ports(443, 50000))), up1))
    P.append((eth(other, SMAC, 0x86DD, ipv6(C6, D6, 17, ports(443, 50000)), tags=[(0x8100, 0x0FFF)]), up1))
    P.append((eth(other, SMAC, 0x86DD, ipv6(C6, D6, 0, ports(443, 50000))), up1))          # extension header
    P.append((eth(other, SMAC, 0x86DD, ipv6(C6, D6, 6, ports(443, 50000), version=4)), up1))
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80)), tags=[(0x8100, 5), (0x8100, 6)]), up0))  # QinQ
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80)), tags=[(0x88A8, 5)]), up0))

    # ---- SFC L2 steering ----------------------------------------------------------------
    arp = bytes.fromhex("0001080006040001") + bytes(20)
    P.append((eth(mac0, SMAC, 0x0806, arp), vf0))              # R1 vf -> nf0.in
    P.append((eth(mac0, SMAC, 0x0806, arp), nf_in[0]))         # R2 nf0.in, dl_dst=mac0 -> vf0
    P.append((eth(mac1, SMAC, 0x0806, arp), nf_in[0]))         # R2 -> vf1
    P.append((eth(other, SMAC, 0x0806, arp), nf_in[0]))        # no rule for this MAC, no wildcard -> miss
    P.append((eth(mac1, SMAC, 0x0806, arp), nf_out[1]))        # R3 hairpin -> in_port
    P.append((eth(other, SMAC, 0x0806, arp), nf_out[1]))       # R1 last.out -> wire
    P.append((eth(other, SMAC, 0x0806, arp), wire))            # R1 wire -> last.out
    P.append((eth(mac0, SMAC, 0x0806, arp), nf_in[1]))         # chain: nf1.in, dl_dst -> nf0.out
    P.append((eth(other, SMAC, 0x0806, arp), nf_out[0]))       # chain: nf0.out -> nf1.in
    P.append((eth(mac0, SMAC, 0x0806, arp), 999))              # unknown port -> miss
    P.append((eth(mac0, SMAC, 0x0800, ipv4(B, B, 6, ports(1, 2))), vf0))  # IP traffic without a flow -> L2

    # ---- LPM ----------------------------------------------------------------------------
    sc.lpm_add(4, bytes([172, 0, 0, 0]), 8, 1)
    sc.lpm_add(4, bytes([172, 16, 0, 0]), 12, 2)
    sc.lpm_add(4, bytes([172, 16, 5, 0]), 24, nf_in[1])
    sc.lpm_add(4, bytes([172, 16, 5, 128]), 25, 4)
    sc.lpm_add(4, bytes([172, 16, 5, 200]), 32, 5)
    sc.lpm_add(4, bytes([192, 168, 255, 255]), 17, 6)  # host bits set: must be ignored
    sc.lpm_add(6, bytes.fromhex("2a000000000000000000000000000000"), 12, 1)
    sc.lpm_add(6, bytes.fromhex("2a001450000000000000000000000000"), 32, 2)
    sc.lpm_add(6, bytes.fromhex("2a001450400100000000000000000000"), 48, 3)
    sc.lpm_add(6, bytes.fromhex("2a001450400108000000000000000000"), 56, 4)
    sc.lpm_add(6, bytes.fromhex("2a0014504001080b0000000000000000"), 64, 5)
    sc.lpm_add(6, bytes.fromhex("2a0014504001080b000000000000200e"), 128, 6)
    sc.lpm_add(6, bytes.fromhex("2a0014504001080b0000000000002000"), 121, 1)
    for last in ([172, 1, 2, 3], [172, 16, 9, 9], [172, 16, 5, 1], [172, 16, 5, 129], [172, 16, 5, 200], [172, 31, 255, 255],
                 [172, 32, 0, 0], [192, 168, 128, 1], [192, 168, 127, 1], [8, 8, 8, 8]):
        P.append((eth(other, SMAC, 0x0800, ipv4(A, bytes(last), 17, ports(5, 6))), up0))
    for d in ("2a0f0000000000000000000000000001", "2a001450000000000000000000000001", "2a001450400100000000000000000001",
              "2a0014504001080000000000000000aa", "2a0014504001080b00000000deadbeef", "2a0014504001080b000000000000200e",
              "2a0014504001080b000000000000207f", "2a0014504001080b0000000000002080", "2b000000000000000000000000000000"):
        P.append((eth(other, SMAC, 0x86DD, ipv6(C6, bytes.fromhex(d), 17, ports(5, 6))), up1))

    # ---- ACL ----------------------------------------------------------------------------
    X = bytes([203, 0, 113, 0])
    sc.acl_add(acl(4, dst=X, dst_len=24, priority=10, deny=True))                    # id 0
    sc.acl_add(acl(4, dst=X[:3] + b"\x07", dst_len=32, priority=20, deny=False))     # id 1: permit beats deny
    sc.acl_add(acl(4, dst=X[:3] + b"\x08", dst_len=32, priority=10, deny=False))     # id 2: tie with id 0 -> id 0 wins
    sc.acl_add(acl(0, dport=23, priority=5, deny=True))                              # id 3: any version, telnet
    sc.acl_add(acl(6, src=C6, src_len=64, proto=17, dport=0x1F00, dport_mask=0xFF00, priority=7, deny=True))  # id 4
    sc.acl_add(acl(4, in_port=up1, vlan=77, priority=3, deny=True))                  # id 5
    sc.acl_add(acl(4, dst=X, dst_len=24, priority=10, deny=True))                    # id 6: duplicate of id 0, loses the tie
    for last, dp_, vl, port in ((1, 80, None, up0), (7, 80, None, up0), (8, 80, None, up0), (1, 23, None, up0)):
        P.append((eth(other, SMAC, 0x0800, ipv4(A, X[:3] + bytes([last]), 6, ports(9, dp_))), port))
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(9, 23))), up0))                       # telnet over v4
    P.append((eth(other, SMAC, 0x86DD, ipv6(C6, D6, 6, ports(9, 23))), up1))                     # telnet over v6
    P.append((eth(other, SMAC, 0x0806, arp[:20] + b"\x00\x17" + bytes(6)), up0))                 # non-IP: ACL skipped
    P.append((eth(other, SMAC, 0x86DD, ipv6(C6, D6, 17, ports(9, 0x1F42))), up1))                # masked dport
    P.append((eth(other, SMAC, 0x86DD, ipv6(C6, D6, 17, ports(9, 0x2042))), up1))
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80)), tags=[(0x8100, 77)]), up1))  # in_port + vlan
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80)), tags=[(0x8100, 77)]), up0))
    P.append((eth(other, SMAC, 0x0800, ipv4(A, B, 6, ports(1000, 80), frag=185) ), up0))

    hdr = np.frombuffer(b"".join(f for f, _ in P), dtype=np.uint8).reshape(-1, 64).copy()
    in_port = np.array([p for _, p in P], dtype=np.uint16)
    return hdr, in_port
