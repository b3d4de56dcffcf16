"""ACL forests under rules the synthetic generator never makes: ARBITRARY (non-prefix, non-contiguous) masks on
every field, priority ties, duplicates, rules on in_port / vlan, IPv6 rules that constrain the low address words,
catch-alls in the middle of the ranking.  The built image (walked in Python) must name the same winner as a
brute-force scan in (priority desc, id asc) order, for every forest shape; the CUDA path must equal the oracle."""
import numpy as np
import pytest

from common import Scenario, consts, sfc
from packets import eth, ipv4, ipv6, ports
from test_image_walk import Image, acl_key

FIELDS16 = ("in_port", "vlan", "sport", "dport")


def random_rules(rng, n):
    rules = np.zeros(n, dtype=sfc.ACL_RULE_DTYPE)
    for i in range(n):
        r = rules[i]
        ver = int(rng.choice([0, 4, 4, 6]))
        r["ip_ver"] = ver
        if ver:
            nbytes = 4 if ver == 4 else 16
            for f in ("src", "dst"):
                kind = int(rng.integers(0, 4))
                if kind == 0:
                    continue  # wildcard
                mask = np.zeros(16, np.uint8)
                if kind == 1:  # prefix
                    ln = int(rng.integers(1, nbytes * 8 + 1))
                    m = ((1 << ln) - 1) << (128 - ln)
                    mask = np.frombuffer(m.to_bytes(16, "big"), np.uint8).copy()
                else:  # arbitrary bits, possibly only in the low words
                    mask[:nbytes] = rng.integers(0, 256, nbytes).astype(np.uint8) & rng.integers(0, 256, nbytes).astype(np.uint8)
                    if kind == 3 and ver == 6:
                        mask[:8] = 0
                mask[nbytes:] = 0
                val = np.zeros(16, np.uint8)
                val[:nbytes] = rng.integers(0, 4, nbytes).astype(np.uint8) * 85  # few distinct byte values: rules overlap
                r[f] = val & mask
                r[f + "_mask"] = mask
        if rng.integers(0, 2):
            r["proto"], r["proto_mask"] = int(rng.choice([6, 17])), int(rng.choice([0xFF, 0xFF, 0x10]))
            r["proto"] &= r["proto_mask"]
        for f in FIELDS16:
            if rng.integers(0, 4) == 0:
                m = int(rng.choice([0xFFFF, 0xFF00, 0x00FF, 0x0F0F, 0x8001]))
                if f == "vlan":
                    m &= 0x0FFF
                r[f + "_mask"] = m
                r[f] = int(rng.integers(0, 4)) * 0x1111 & m
        r["priority"] = int(rng.integers(1, 6))  # heavy ties: order falls back on the id
        r["verdict"] = consts.ACL_DENY if rng.integers(0, 3) == 0 else consts.ACL_PERMIT
    rules[n // 2] = np.zeros(1, dtype=sfc.ACL_RULE_DTYPE)[0]  # a catch-all (any version) in the middle of the ranking
    rules[n // 2]["priority"] = 3
    rules[n // 2 + 1] = rules[n // 3]  # an exact duplicate of another rule
    return rules


def random_packets(rng, n):
    frames, in_ports = [], []
    mac = bytes(6)
    for _ in range(n):
        v6 = rng.integers(0, 3) == 0
        nb = 16 if v6 else 4
        src = bytes((rng.integers(0, 4, nb) * 85).astype(np.uint8))
        dst = bytes((rng.integers(0, 4, nb) * 85).astype(np.uint8))
        l4 = ports(int(rng.integers(0, 4)) * 0x1111, int(rng.integers(0, 4)) * 0x1111)
        proto = int(rng.choice([6, 17, 1]))
        ip = ipv6(src, dst, proto, l4) if v6 else ipv4(src, dst, proto, l4)
        tags = [(0x8100, int(rng.integers(0, 4)) * 0x111)] if rng.integers(0, 2) else []
        frames.append(np.frombuffer(eth(mac, mac, 0x86DD if v6 else 0x0800, ip, tags=tags), np.uint8))
        in_ports.append(int(rng.integers(0, 4)) * 0x1111 & 0x0FFF)
    return np.stack(frames), np.array(in_ports, np.uint16)


def brute(rules, ids, q):
    order = sorted(range(len(rules)), key=lambda i: (-int(rules[i]["priority"]), ids[i]))
    for i in order:
        r = rules[i]
        w = lambda b: int.from_bytes(bytes(b), "big")  # noqa: E731
        val = [(int(r["in_port"]) << 16) | int(r["vlan"]), (int(r["ip_ver"]) << 24) | (int(r["proto"]) << 16) | int(r["sport"]), int(r["dport"]) << 16]
        msk = [(int(r["in_port_mask"]) << 16) | int(r["vlan_mask"]), ((0xFF if r["ip_ver"] else 0) << 24) | (int(r["proto_mask"]) << 16) | int(r["sport_mask"]),
               int(r["dport_mask"]) << 16]
        for f in ("src", "dst"):
            val += [w(r[f][j:j + 4]) for j in range(0, 16, 4)]
            msk += [w(r[f + "_mask"][j:j + 4]) for j in range(0, 16, 4)]
        if all(((q[j] ^ val[j]) & msk[j]) == 0 for j in range(11)):
            return ids[i], int(r["verdict"]) == consts.ACL_DENY
    return None


@pytest.mark.parametrize("feat", ["", "0", "5", "7"])
@pytest.mark.parametrize("bits", ["", "4"])
def test_forest_names_the_brute_force_winner(feat, bits, monkeypatch):
    if feat:
        monkeypatch.setenv("SFC_ACL_FEAT", feat)
    if bits:
        monkeypatch.setenv("SFC_ACL_BITS", bits)
    rng = np.random.default_rng(17)
    rules = random_rules(rng, 160)
    sc = Scenario(sfc.Dataplane(-1))
    ids = [sc.acl_add(rules[i:i + 1]) for i in range(len(rules))]
    sc.dp.commit()
    img = Image(sc.dp.table_export())
    hdr, in_port = random_packets(rng, 700)
    matched = denied = 0
    for i in range(len(in_port)):
        p, q = acl_key(hdr[i], int(in_port[i]))
        if p["ver"] == 0:
            continue
        got, want = img.acl(q), brute(rules, ids, q)
        if want is None:
            assert got is None, i
        else:
            assert got is not None and (got & 0x0FFFFFFF, got >> 31 == 1) == want, (i, want, got)
            matched += 1
            denied += want[1]
    assert matched > 300 and denied > 20


@pytest.mark.gpu
def test_cuda_path_on_arbitrary_masks(torch_cuda):
    rng = np.random.default_rng(23)
    rules = random_rules(rng, 300)
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    for i in range(len(rules)):
        sc.acl_add(rules[i:i + 1])
    dp.commit()
    hdr, in_port = random_packets(rng, 20000)
    t = torch_cuda
    d_hdr, d_port = t.from_numpy(hdr).cuda(), t.from_numpy(in_port.view(np.int16)).cuda()
    res = t.zeros(len(in_port), dtype=t.int64, device="cuda")
    dp.classify(d_hdr.data_ptr(), d_port.data_ptr(), res.data_ptr(), len(in_port), consts.F_COUNTERS, t.cuda.current_stream().cuda_stream)
    t.cuda.synchronize()
    got = res.cpu().numpy().view(np.uint64)
    want, cnt = sc.orc.classify(hdr, in_port, counters=True)
    assert np.array_equal(got, want)
    assert np.array_equal(dp.counters_read(), cnt)
    assert (sfc.result_fields(got)[3] == consts.STAGE_ACL).sum() > 20  # denies; permits show in the per-rule counters
    assert (cnt[consts.CNT_RULES0:consts.CNT_RULES0 + 300] > 0).sum() > 2
