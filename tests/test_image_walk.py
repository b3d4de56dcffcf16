"""CPU checks of the table-image BUILDER (no GPU): a small Python walker reads the exported image
with the documented layout (csrc/sfc_common.h) and follows the same structures the kernel follows —
the ACL decision forest, the LPM tries and the cuckoo buckets — and its answers must equal the
oracle's.  The walker lives in the tests on purpose: the library has no CPU lookup path."""
import struct

import numpy as np
import pytest

from common import Oracle, Scenario, config3_spec, config4_spec, consts, install_flows, install_lpm_acl, sfc

M64 = (1 << 64) - 1
GAMMA = 0x9E3779B97F4A7C15


def mix64(z):
    z &= M64
    z ^= z >> 30
    z = (z * 0xBF58476D1CE4E5B9) & M64
    z ^= z >> 27
    z = (z * 0x94D049BB133111EB) & M64
    return z ^ (z >> 31)


def expand_seed(seed, n):
    return [((mix64(seed + (i + 1) * GAMMA) >> 16) & 0xFFFFFFFF) | 1 for i in range(n)]


class Image:
    """Field offsets follow struct ImageHeader (version 12)."""

    def __init__(self, img: np.ndarray):
        self.b = img.tobytes()
        u = lambda fmt, off: struct.unpack_from("<" + fmt, self.b, off)  # noqa: E731
        magic, ver, hb, total, epoch = u("QIIQQ", 0)
        assert magic == 0x3130434653303042 and ver == 12 and hb == 512 and total == len(self.b)
        o = 32
        self.f4_off, self.f4_seed, self.f4_nb, self.f4_nh, self.f4_cnt, self.f4_first = u("QQIIII", o); o += 32
        self.f6_off, self.f6_seed, self.f6_nb, self.f6_nh, self.f6_cnt, self.f6_first = u("QQIIII", o); o += 32
        self.l2x_off, self.l2x_mask, self.l2x_cnt, self.l2w_off, self.l2w_n, self.l2w_cnt = u("QIIQII", o); o += 32
        self.lpm4 = u("QQQII", o); o += 32
        self.lpm6 = u("QQQII", o); o += 32
        self.acl_nodes_off, self.acl_leaf_off, self.acl_rule_off, self.acl_res_off = u("QQQQ", o); o += 32
        self.acl_nrules, self.acl_nentries, self.acl_nleaf, self.acl_ntrees = u("IIII", o); o += 16
        roots = u("32I", o); o += 128
        self.acl_roots = [roots[4 * i: 4 * i + 4] for i in range(8)]
        self.lpm4_tail_off, self.lpm6_tail_off, self.lpm4_ntails, self.lpm6_ntails = u("QQII", o); o += 24
        self.f4_slots, self.f6_slots = u("II", o); o += 8
        self.flow_format, self.fidx_bits = u("II", o); o += 8
        self.fidx_off, self.fidx_seed, self.frec_off = u("QQQ", o); o += 24
        self.fidx_nb, self.frec_units, self.frec_used, self.fidx_first = u("IIII", o); o += 16
        self.l2x_copy_off, self.l2w_copy_off = u("QQ", o), u("QQ", o + 16); o += 32
        self.l2x_cap, self.l2_active = u("II", o); o += 8
        assert o == 488 and self.l2x_off == self.l2x_copy_off[self.l2_active] and self.l2w_off == self.l2w_copy_off[self.l2_active]

    def u32(self, off, n):
        return np.frombuffer(self.b, dtype="<u4", count=n, offset=off)

    # ---- ACL forest: returns winning rule word (id | deny<<31) or None ----
    def acl(self, q, stats=None):  # q: 11 key words
        LEAF, MORE, OFF, NONE = 1 << 31, 1 << 30, (1 << 28) - 1, 0xFFFF
        groups = ((1, 2, 7, 3), (0, 8, 9, 10), (4, 5, 6))
        rows = list(q) + [0]  # row 11 is the zero row unused fields point at
        best = None

        def matches(c):
            if stats is not None:
                stats["compares"] += 1
            rec = self.u32(self.acl_rule_off + (c & 0x0FFF) * 96, 24)
            ok = all(((q[w] ^ int(rec[j])) & int(rec[4 + j])) == 0 for j, w in enumerate(groups[0]))
            if ok and c & 0x1000:
                ok = all(((q[w] ^ int(rec[8 + j])) & int(rec[12 + j])) == 0 for j, w in enumerate(groups[1]))
            if ok and c & 0x2000:
                ok = all(((q[w] ^ int(rec[16 + j])) & int(rec[20 + j])) == 0 for j, w in enumerate(groups[2]))
            return ok

        for t in range(self.acl_ntrees):
            e = self.acl_roots[t]
            while not e[0] & LEAF:
                idx = 0
                for f in (e[1] & 0xFFFF, e[1] >> 16, e[2] & 0xFFFF, e[2] >> 16, e[3] & 0xFFFF, e[3] >> 16):
                    assert f & 0x60 == 0 and (f >> 7) <= 11
                    idx = (idx << 1) | ((rows[f >> 7] >> (f & 31)) & 1)
                assert e[0] + idx < self.acl_nentries
                e = tuple(int(x) for x in self.u32(self.acl_nodes_off + (e[0] + idx) * 16, 4))
                if stats is not None:
                    stats["visits"] += 1
            cands = [e[1] & 0xFFFF, e[1] >> 16, e[2] & 0xFFFF, e[2] >> 16, e[3] & 0xFFFF, e[3] >> 16]
            if e[0] & MORE:
                assert NONE not in cands
                off = self.acl_leaf_off + (e[0] & OFF) * 2
                cnt = int(np.frombuffer(self.b, dtype="<u2", offset=off, count=1)[0])
                cands += [int(c) for c in np.frombuffer(self.b, dtype="<u2", offset=off + 2, count=cnt)]
            for c in cands:
                if c == NONE or (best is not None and (c & 0x0FFF) >= best):
                    break
                if matches(c):
                    best = c & 0x0FFF
                    break
        return None if best is None else int(self.u32(self.acl_res_off + best * 4, 1)[0])

    def lpm(self, af, addr_words):
        root_off, nodes_off, nh_off, nnodes, cnt = self.lpm4 if af == 4 else self.lpm6
        e = int(self.u32(root_off + (addr_words[0] >> 16) * 4, 1)[0])
        pos = 16
        while e & 0x80000000 and pos < (32 if af == 4 else 128):
            byte = (addr_words[pos >> 5] >> (24 - (pos & 31))) & 0xFF
            e = int(self.u32(nodes_off + ((e & 0x7FFFFFFF) * 256 + byte) * 4, 1)[0])
            pos += 8
        if e & 0x80000000:
            return None
        if e & 0x40000000:  # a lone prefix: one record, one masked compare
            rec = self.u32((self.lpm4_tail_off if af == 4 else self.lpm6_tail_off) + (e & 0x3FFFFFFF) * 32, 8)
            ln, diff = int(rec[4]), 0
            for i in range(4):
                bits = min(32, max(0, ln - 32 * i))
                mask = (0xFFFFFFFF << (32 - bits)) & 0xFFFFFFFF if bits else 0
                diff |= (addr_words[i] ^ int(rec[i])) & mask
            e = int(rec[6]) if diff else int(rec[5])
            if not diff:  # a tail hit carries the high result word itself
                hi, lo = self.u32(nh_off + (e - 1) * 8, 2)
                assert int(rec[7]) == int(hi) and int(lo) == ((consts.STAGE_LPM4 if af == 4 else consts.STAGE_LPM6) << 28 | (e - 1))
        if e == 0:
            return None
        hi, lo = self.u32(nh_off + (e - 1) * 8, 2)
        return (int(hi) << 32) | int(lo)

    def flow4(self, sip, dip, ports, meta20):
        S = expand_seed(self.f4_seed, 10)
        P = ((sip ^ S[0]) * (dip ^ S[1])) & M64
        Q = ((ports ^ S[2]) * (meta20 ^ S[3])) & M64
        x = (P & 0xFFFFFFFF) ^ (Q >> 32)
        y = (P >> 32) ^ (Q & 0xFFFFFFFF)
        for j in range(self.f4_nh):
            F = ((x ^ S[4 + 2 * j]) * (y ^ S[5 + 2 * j])) & M64
            h = (F & 0xFFFFFFFF) ^ (F >> 32)
            b = (h * self.f4_nb) >> 32
            for sl in range(self.f4_slots):
                s = self.u32(self.f4_off + (b * self.f4_slots + sl) * 32, 8)
                if (int(s[0]), int(s[1]), int(s[2]), int(s[3])) == (sip, dip, ports, meta20 | 0x80000000):
                    return (int(s[4]) << 32) | int(s[5]), j
        return None, None


def _flow6(self, sip, dip, ports, meta20):
    """sip / dip: four words each.  Mirrors flow6_hash + flow_bucket + the 64-byte slot compare."""
    S = expand_seed(self.f6_seed, 16)
    mw = lambda a, b: (a * b) & M64  # noqa: E731
    P1, P2 = mw(sip[0] ^ S[0], sip[1] ^ S[1]), mw(sip[2] ^ S[2], sip[3] ^ S[3])
    P3, P4 = mw(dip[0] ^ S[4], dip[1] ^ S[5]), mw(dip[2] ^ S[6], dip[3] ^ S[7])
    Q = mw(ports ^ S[8], meta20 ^ S[9])
    lo, hi = (lambda v: v & 0xFFFFFFFF), (lambda v: v >> 32)
    x = lo(P1) ^ hi(P2) ^ lo(P3) ^ hi(P4) ^ hi(Q)
    y = hi(P1) ^ lo(P2) ^ hi(P3) ^ lo(P4) ^ lo(Q)
    for j in range(self.f6_nh):
        F = mw(x ^ S[10 + 2 * j], y ^ S[11 + 2 * j])
        b = ((lo(F) ^ hi(F)) * self.f6_nb) >> 32
        for sl in range(self.f6_slots):
            s = [int(v) for v in self.u32(self.f6_off + (b * self.f6_slots + sl) * 64, 12)]
            if s[:4] == list(sip) and s[4:8] == list(dip) and s[8] == ports and s[9] == meta20 | 0x80000000:
                return (s[10] << 32) | s[11], j
    return None, None


def _flow2(self, ver, sip, dip, ports, meta20):
    """Format 2 (sfc_common.h): flow6_hash over the full key -> two index buckets of 7 entries + overflow count ->
    fingerprint match -> record unit(s).  Returns (result word, which bucket) or (None, None)."""
    S = expand_seed(self.fidx_seed, 16)
    v6 = ver == 6
    sip, dip = list(sip) + [0] * (4 - len(sip)), list(dip) + [0] * (4 - len(dip))
    mw = lambda a, b: (a * b) & M64  # noqa: E731
    lo, hi = (lambda v: v & 0xFFFFFFFF), (lambda v: v >> 32)
    meta = meta20 | (0x100000 if v6 else 0)
    P1, P2 = mw(sip[0] ^ S[0], sip[1] ^ S[1]), mw(sip[2] ^ S[2], sip[3] ^ S[3])
    P3, P4 = mw(dip[0] ^ S[4], dip[1] ^ S[5]), mw(dip[2] ^ S[6], dip[3] ^ S[7])
    Q = mw(ports ^ S[8], meta ^ S[9])
    x = lo(P1) ^ hi(P2) ^ lo(P3) ^ hi(P4) ^ hi(Q)
    y = hi(P1) ^ lo(P2) ^ hi(P3) ^ lo(P4) ^ lo(Q)
    F = mw(x ^ S[14], y ^ S[15])
    B = self.fidx_bits
    fp, mask = (lo(F) ^ hi(F)) >> B, (1 << B) - 1
    tag = meta20 | 0x80000000 | (0x40000000 if v6 else 0)
    buckets = []
    for j in range(2):
        G = mw(x ^ S[10 + 2 * j], y ^ S[11 + 2 * j])
        buckets.append(((lo(G) ^ hi(G)) * self.fidx_nb) >> 32)
    for j, b in enumerate(buckets):
        if j == 1 and (b == buckets[0] or overflow == 0):  # noqa: F821  (set in the first round)
            break
        e = [int(v) for v in self.u32(self.fidx_off + b * 32, 8)]
        overflow = e[7]
        for ent in e[:7]:
            if ent == 0 or ent >> B != fp:
                continue
            unit = (ent & mask) - 1
            assert 0 <= unit < self.frec_units
            u = [int(v) for v in self.u32(self.frec_off + unit * 32, 16)]
            want = [tag, ports] + ([sip[0], dip[0], 0, 0] if not v6 else sip)
            if [u[0], u[1]] + u[4:8] == want and (not v6 or u[8:12] == dip):
                return (u[2] << 32) | u[3], j
    return None, None


def _flow(self, k):
    """Look a flow key (FLOW_KEY_DTYPE record) up in whichever flow table format the image uses."""
    ver = int(k["ip_ver"])
    w4 = lambda b: [int.from_bytes(bytes(b[i:i + 4]), "big") for i in range(0, 16, 4)]  # noqa: E731
    ports, meta20 = (int(k["sport"]) << 16) | int(k["dport"]), int(k["proto"]) | (int(k["vlan"]) << 8)
    if self.flow_format == 2:
        return self.flow2(ver, w4(k["src"]), w4(k["dst"]), ports, meta20)
    if ver == 4:
        return self.flow4(w4(k["src"])[0], w4(k["dst"])[0], ports, meta20)
    return self.flow6(w4(k["src"]), w4(k["dst"]), ports, meta20)


def _l2(self, in_port, mac):
    """Exact (in_port, dl_dst) first, then the in_port wildcard; returns the result word or None."""
    key_hi, key_lo = (in_port << 16) | int.from_bytes(mac[:2], "big"), int.from_bytes(mac[2:], "big")
    if self.l2x_cnt:
        P = ((key_hi ^ 0x9E3779B1) * (key_lo ^ 0x85EBCA6B)) & M64
        i = ((P & 0xFFFFFFFF) ^ (P >> 32)) & self.l2x_mask
        while True:
            e = [int(v) for v in self.u32(self.l2x_off + i * 16, 4)]
            if not e[3]:
                break
            if e[0] == key_lo and e[1] == key_hi:
                return (e[3] << 32) | e[2]
            i = (i + 1) & self.l2x_mask
    if self.l2w_cnt and in_port < self.l2w_n:
        hi, lo = (int(v) for v in self.u32(self.l2w_off + in_port * 8, 2))
        if hi:
            return (hi << 32) | lo
    return None


Image.flow6 = _flow6
Image.flow2 = _flow2
Image.flow = _flow
Image.l2 = _l2


def acl_key(pkt, in_port):
    p = Oracle.parse(pkt)
    w = lambda b: int.from_bytes(b, "big")  # noqa: E731
    q = [(in_port << 16) | p["vlan"], (p["ver"] << 24) | (p["proto"] << 16) | p["sport"], p["dport"] << 16]
    q += [w(p["src"][i:i + 4]) for i in range(0, 16, 4)] + [w(p["dst"][i:i + 4]) for i in range(0, 16, 4)]
    return p, q


def test_acl_forest_and_lpm_tries_answer_like_the_oracle():
    sc = Scenario(sfc.Dataplane(-1))
    spec = config4_spec(n_pfx4=3000, n_pfx6=800, n_acl=400)
    rules = install_lpm_acl(sc, spec)
    sc.dp.commit()
    img = Image(sc.dp.table_export())
    assert img.acl_nrules == 400 and 1 <= img.acl_ntrees <= 8
    hdr, ports = sfc.gen_packets_host(spec, 0, 3000)
    want = sc.orc.classify(hdr, ports)
    act, _, _, stage, index = sfc.result_fields(want)
    checked_acl = checked_lpm = 0
    for i in range(len(ports)):
        p, q = acl_key(hdr[i], int(ports[i]))
        if p["ver"] == 0:
            continue
        r = img.acl(q)
        if stage[i] == consts.STAGE_ACL:  # the oracle dropped it: the forest must name the same deny rule
            assert r is not None and r >> 31 == 1 and (r & 0x0FFFFFFF) == index[i], i
            checked_acl += 1
            continue
        assert r is None or r >> 31 == 0, f"packet {i}: forest says deny, oracle let it pass"
        dst = [q[7], q[8], q[9], q[10]]
        got = img.lpm(p["ver"], dst)
        if stage[i] in (consts.STAGE_LPM4, consts.STAGE_LPM6):
            assert got == int(want[i]), i
            checked_lpm += 1
        else:
            assert got is None
    assert checked_acl > 5 and checked_lpm > 1000


def test_acl_forest_on_the_handmade_rules():
    from packets import handmade_cases

    sc = Scenario(sfc.Dataplane(-1))
    hdr, ports = handmade_cases(sc)
    sc.dp.commit()
    img = Image(sc.dp.table_export())
    want = sc.orc.classify(hdr, ports)
    stage, index = sfc.result_fields(want)[3:5]
    for i in range(len(ports)):
        p, q = acl_key(hdr[i], int(ports[i]))
        if p["ver"] == 0:
            continue
        r = img.acl(q)
        if stage[i] == consts.STAGE_ACL:
            assert r is not None and r >> 31 == 1 and (r & 0x0FFFFFFF) == index[i]
        else:
            assert r is None or r >> 31 == 0


@pytest.mark.parametrize("slots", [1, 2])
def test_every_ipv4_flow_sits_in_one_of_its_buckets(slots):
    sc = Scenario(sfc.Dataplane(-1))
    sc.dp.set_option("flow_format", 1)
    sc.dp.set_option("flow_slots", slots)
    spec = config3_spec(n_flows=20000, v6_permille=0, n_ports=8)
    install_flows(sc, spec)
    sc.dp.commit()
    img = Image(sc.dp.table_export())
    assert img.f4_cnt == 20000 and img.f4_slots == slots
    assert img.f4_nb >= 20000 / 0.45 if slots == 1 else 20000 / 2 / 0.86 <= img.f4_nb <= 20000 / 2 / 0.7
    keys, acts = sfc.gen_flows(spec, 0, 20000)
    first = 0
    for i in range(0, 20000, 7):
        k = keys[i]
        sip, dip = int.from_bytes(bytes(k["src"][:4]), "big"), int.from_bytes(bytes(k["dst"][:4]), "big")
        res, j = img.flow4(sip, dip, (int(k["sport"]) << 16) | int(k["dport"]), int(k["proto"]) | (int(k["vlan"]) << 8))
        assert res is not None, f"flow {i} is not where the hash says"
        assert res & 0x0FFFFFFF == i and (res >> 28) & 0xF == consts.STAGE_FLOW and res >> 56 == int(acts[i]["action"])
        first += j == 0
    assert first / len(range(0, 20000, 7)) > 0.6  # most flows answer in their first bucket
    assert abs(img.f4_first / 20000 - first / len(range(0, 20000, 7))) < 0.05  # header statistic is honest


def test_indexed_format_finds_every_flow_of_both_families():
    """Format 2 (option flow_format): every flow is found through the fingerprint index, nearly all in their
    first bucket, and keys that were never added are not found."""
    sc = Scenario(sfc.Dataplane(-1))
    sc.dp.set_option("flow_format", 2)
    spec = config3_spec(n_flows=30000, v6_permille=300, n_ports=8)
    install_flows(sc, spec)
    sc.dp.commit()
    img = Image(sc.dp.table_export())
    keys, acts = sfc.gen_flows(spec, 0, 31000)
    n6 = int((keys[:30000]["ip_ver"] == 6).sum())
    assert img.flow_format == 2 and img.f4_cnt + img.f6_cnt == 30000 and img.f6_cnt == n6
    assert img.frec_used == 30000 + n6 and img.fidx_nb * 7 * 0.6 <= 30000 <= img.fidx_nb * 7 * 0.7
    first = 0
    for i in range(0, 30000, 7):
        res, j = img.flow(keys[i])
        assert res is not None, f"flow {i} is not where the hash says"
        assert res & 0x0FFFFFFF == i and (res >> 28) & 0xF == consts.STAGE_FLOW and res >> 56 == int(acts[i]["action"])
        first += j == 0
    assert first / len(range(0, 30000, 7)) > 0.9
    assert abs(img.fidx_first / 30000 - first / len(range(0, 30000, 7))) < 0.03
    for i in range(30000, 31000):
        assert img.flow(keys[i])[0] is None


def test_forced_forests_answer_like_the_single_tree(monkeypatch):
    """The category split (SFC_ACL_FEAT) changes the shape of the ACL structure, never its answers."""
    spec = config4_spec(n_pfx4=500, n_pfx6=100, n_acl=600)
    hdr, ports = sfc.gen_packets_host(spec, 0, 1500)
    answers = []
    for feat in ("0", "1", "5", "7"):
        monkeypatch.setenv("SFC_ACL_FEAT", feat)
        sc = Scenario(sfc.Dataplane(-1))
        install_lpm_acl(sc, spec)
        sc.dp.commit()
        img = Image(sc.dp.table_export())
        out = []
        for i in range(len(ports)):
            p, q = acl_key(hdr[i], int(ports[i]))
            out.append(None if p["ver"] == 0 else img.acl(q))
        answers.append((img.acl_ntrees, out))
    assert answers[0][0] == 1 and answers[3][0] > answers[1][0] >= 2
    assert all(a[1] == answers[0][1] for a in answers[1:])


def test_import_refuses_images_whose_indices_point_outside_their_tables():
    """sfc_table_import follows every index the kernel would follow (LPM children, tails and next hops,
    ACL child arrays, candidates and overflow lists) and refuses the image instead of letting the device fault."""
    import errno
    import struct as st
    sc = Scenario(sfc.Dataplane(-1))
    spec = config4_spec(n_pfx4=3000, n_pfx6=800, n_acl=400)
    install_lpm_acl(sc, spec)
    sc.dp.commit()
    good = sc.dp.table_export()
    img = Image(good)
    other = sfc.Dataplane(-1)
    other.table_import(good)  # the untouched image is accepted
    assert np.array_equal(other.table_export(), good)

    def refused(mutate):
        bad = good.copy()
        mutate(bad)
        with pytest.raises(sfc.SfcError) as e:
            sfc.Dataplane(-1).table_import(bad)
        assert e.value.code == -errno.EINVAL
        return str(e.value)

    def put32(buf, off, v):
        buf[off:off + 4] = np.frombuffer(st.pack("<I", v), dtype=np.uint8)

    root4 = img.lpm4[0]
    assert "LPM" in refused(lambda b: put32(b, root4 + 4 * 77, 0x80000000 | img.lpm4[3]))       # child past the last node
    assert "LPM" in refused(lambda b: put32(b, root4 + 4 * 77, 0x40000000 | img.lpm4_ntails))   # tail past the last record
    assert "LPM" in refused(lambda b: put32(b, root4 + 4 * 77, 0x3FFFFFFF))                     # next hop far outside
    assert img.lpm6_ntails > 0
    assert "LPM" in refused(lambda b: put32(b, img.lpm6_tail_off + 16, 129))                    # tail length > 128
    assert "ACL" in refused(lambda b: put32(b, img.acl_nodes_off, img.acl_nentries))            # child array past the end
    leaf = next(i for i in range(img.acl_nentries) if int(img.u32(img.acl_nodes_off + 16 * i, 1)[0]) >> 31)
    assert "ACL" in refused(lambda b: put32(b, img.acl_nodes_off + 16 * leaf + 4, 0x0FFF))      # candidate rank >= nrules
    assert "ACL" in refused(lambda b: put32(b, img.acl_nodes_off + 16 * leaf, 0xC0000000 | img.acl_nleaf))  # overflow list outside
    assert "ACL" in refused(lambda b: put32(b, 32 * 6 + 32 + 16, img.acl_nentries))             # root entry (header)
    assert refused(lambda b: put32(b, 8, 8))                                                     # older image version
    with pytest.raises(sfc.SfcError):
        sfc.Dataplane(-1).table_import(good[:-256])                                               # truncated


@pytest.mark.parametrize("fmt", [1, 2])
def test_ipv6_flows_and_l2_rules_are_where_the_kernel_will_look(fmt):
    """The two table kinds the other walks do not reach: every IPv6 flow in one of its buckets with the result the
    API was given, and the chain's L2 steering rules answering like the oracle (exact beats wildcard)."""
    sc = Scenario(sfc.Dataplane(-1))
    sc.dp.set_option("flow_format", fmt)
    topo = sc.build_chain(8, 3)
    spec = config3_spec(n_flows=6000, v6_permille=1000, n_ports=topo["n_ports"])
    install_flows(sc, spec)
    sc.dp.commit()
    img = Image(sc.dp.table_export())
    assert img.f6_cnt == 6000 and img.f4_cnt == 0
    keys, acts = sfc.gen_flows(spec, 0, 6000)
    for i in range(0, 6000, 5):
        res, _ = img.flow(keys[i])
        assert res is not None and res & 0x0FFFFFFF == i and (res >> 28) & 0xF == consts.STAGE_FLOW and res >> 56 == int(acts[i]["action"]), i
    # L2: non-IP frames from every port towards every VF MAC and towards an unknown MAC
    import packets
    macs = [sfc.gen_mac(v) for v in range(8)] + [bytes.fromhex("0200deadbeef")]
    frames, in_ports = [], []
    for port in range(topo["n_ports"]):
        for m in macs:
            frames.append(np.frombuffer(packets.eth(m, bytes(6), 0x0806, bytes(28)), np.uint8))
            in_ports.append(port)
    want = sc.orc.classify(np.stack(frames), np.array(in_ports, np.uint16))
    hits = 0
    for f, port, wv in zip(frames, in_ports, want):
        got = img.l2(port, bytes(f[:6]))
        if got is None:
            assert (int(wv) >> 28) & 0xF == consts.STAGE_MISS
        else:
            assert got == int(wv)
            hits += 1
    assert hits > len(frames) // 3
