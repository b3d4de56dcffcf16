"""LPM tries under adversarial nesting: clustered random prefixes of every length from /0 up, lone prefixes
(tail records) under covering shorter ones, addresses aimed at prefix boundaries.  The built image (walked in
Python) and the CUDA path must agree with a brute-force longest-prefix match and with the oracle."""
import numpy as np
import pytest

from common import Scenario, consts, sfc
from packets import eth, ipv4, ipv6, ports
from test_image_walk import Image


def make_prefixes(rng, af, n):
    """Clusters: a few random stems, each extended to many lengths, plus default routes and host routes."""
    bits = 32 if af == 4 else 128
    out = {}
    stems = [int(rng.integers(0, 1 << 62)) << (bits - 62) if bits > 62 else int(rng.integers(0, 1 << 32)) for _ in range(6)]
    while len(out) < n:
        stem = stems[int(rng.integers(0, len(stems)))]
        ln = int(rng.choice([0, 1, 7, 8, 9, 15, 16, 17, 23, 24, 25, 31, 32] + ([33, 47, 48, 56, 63, 64, 65, 96, 120, 127, 128] if af == 6 else [])))
        flip = int(rng.integers(0, 1 << 12)) << max(0, bits - ln - 6) if ln > 12 else 0  # siblings near the stem
        v = (stem ^ flip) & (((1 << ln) - 1) << (bits - ln)) if ln else 0
        out[(v, ln)] = None
    return list(out)


def brute(prefixes, ports_of, addr, bits):
    best = None
    for i, (v, ln) in enumerate(prefixes):
        if ln == 0 or (addr >> (bits - ln)) == (v >> (bits - ln)):
            if best is None or ln > prefixes[best][1]:
                best = i
    return best


def build(af, seed, n=400):
    rng = np.random.default_rng(seed)
    bits = 32 if af == 4 else 128
    prefixes = make_prefixes(rng, af, n)
    sc = Scenario(sfc.Dataplane(-1))
    ids = [sc.lpm_add(af, v.to_bytes(bits // 8, "big"), ln, 1 + i % 7) for i, (v, ln) in enumerate(prefixes)]
    sc.dp.commit()
    addrs = []
    for v, ln in prefixes:  # first and last address of each prefix, and the addresses just outside
        span = 1 << (bits - ln)
        for a in (v, v + span - 1, v - 1, v + span, v ^ (span >> 1 if span > 1 else 1)):
            addrs.append(a % (1 << bits))
    addrs += [int(rng.integers(0, 1 << 62)) << (bits - 62) if bits > 62 else int(rng.integers(0, 1 << 32)) for _ in range(300)]
    return sc, prefixes, ids, addrs, bits


@pytest.mark.parametrize("af,seed", [(4, 1), (4, 2), (6, 3), (6, 4)])
def test_image_walk_equals_brute_force(af, seed):
    sc, prefixes, ids, addrs, bits = build(af, seed)
    img = Image(sc.dp.table_export())
    assert (img.lpm4_ntails if af == 4 else img.lpm6_ntails) > 0, "the set should exercise tail records"
    stage = consts.STAGE_LPM4 if af == 4 else consts.STAGE_LPM6
    for a in addrs:
        words = [(a >> (bits - 32 * (i + 1))) & 0xFFFFFFFF if bits >= 32 * (i + 1) else 0 for i in range(4)]
        got = img.lpm(af, words)
        want = brute(prefixes, None, a, bits)
        if want is None:
            assert got is None, hex(a)
        else:
            assert got is not None and got & 0x0FFFFFFF == ids[want] and (got >> 28) & 0xF == stage, (hex(a), prefixes[want])
            assert (got >> 32) & 0xFFFF == 1 + want % 7


@pytest.mark.gpu
@pytest.mark.parametrize("af,seed", [(4, 11), (6, 12)])
def test_cuda_path_on_nested_prefixes(torch_cuda, af, seed):
    rng = np.random.default_rng(seed)
    bits = 32 if af == 4 else 128
    prefixes = make_prefixes(rng, af, 600)
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    for i, (v, ln) in enumerate(prefixes):
        sc.lpm_add(af, v.to_bytes(bits // 8, "big"), ln, 1 + i % 7)
    dp.commit()
    addrs = []
    for v, ln in prefixes:
        span = 1 << (bits - ln)
        addrs += [x % (1 << bits) for x in (v, v + span - 1, v - 1, v + span)]
    mac = bytes(6)
    frames = []
    for a in addrs:
        dst = a.to_bytes(bits // 8, "big")
        ip = ipv4(bytes([1, 2, 3, 4]), dst, 6, ports(1, 2)) if af == 4 else ipv6(bytes(16), dst, 17, ports(3, 4))
        frames.append(np.frombuffer(eth(mac, mac, 0x0800 if af == 4 else 0x86DD, ip), np.uint8))
    hdr = np.stack(frames)
    in_port = np.full(len(frames), 9, np.uint16)
    t = torch_cuda
    d_hdr, d_port = t.from_numpy(hdr).cuda(), t.from_numpy(in_port.view(np.int16)).cuda()
    res = t.zeros(len(frames), dtype=t.int64, device="cuda")
    dp.classify(d_hdr.data_ptr(), d_port.data_ptr(), res.data_ptr(), len(frames), 0, t.cuda.current_stream().cuda_stream)
    t.cuda.synchronize()
    got = res.cpu().numpy().view(np.uint64)
    want = sc.orc.classify(hdr, in_port)
    assert np.array_equal(got, want)
    stages = set(sfc.result_fields(got)[3].tolist())
    assert (consts.STAGE_LPM4 if af == 4 else consts.STAGE_LPM6) in stages
