"""GPU parity of sfc_classify_batches: many independent batches in one launch must give, batch by
batch, exactly what one sfc_classify per batch gives, which in turn is the oracle's answer."""
import errno

import numpy as np
import pytest

from common import Scenario, config3_spec, consts, install_flows, install_lpm_acl, sfc
from test_gpu_parity import assert_same

pytestmark = pytest.mark.gpu


class Pool:
    """Batches carved out of separate device buffers at chosen 64-byte offsets."""

    def __init__(self, torch, spec):
        self.torch, self.spec = torch, spec
        self.keep, self.batches, self.hdr_np, self.port_np = [], [], [], []
        self.next_pkt = 0

    def add(self, n, skew=0):
        t = self.torch
        hdr_h, port_h = sfc.gen_packets_host(self.spec, self.next_pkt, n) if n else (np.zeros((0, 64), np.uint8), np.zeros(0, np.uint16))
        self.next_pkt += n
        hdr = t.zeros(n * 64 + skew * 64 + 64, dtype=t.uint8, device="cuda")
        ports = t.zeros(n + 1, dtype=t.int16, device="cuda")
        res = t.full((n + 1,), -1, dtype=t.int64, device="cuda")
        if n:
            hdr[skew * 64: skew * 64 + n * 64] = t.from_numpy(np.array(hdr_h, dtype=np.uint8).reshape(-1)).cuda()
            ports[:n] = t.from_numpy(np.array(port_h, dtype=np.uint16).view(np.int16)).cuda()
        self.keep.append((hdr, ports, res))
        self.batches.append((hdr.data_ptr() + skew * 64, ports.data_ptr(), res.data_ptr(), n))
        self.hdr_np.append(np.asarray(hdr_h).reshape(n, 64))
        self.port_np.append(np.asarray(port_h))

    def results(self):
        self.torch.cuda.synchronize()
        out = []
        for (_, _, res), (_, _, _, n) in zip(self.keep, self.batches):
            r = res.cpu().numpy()
            assert r[n] == -1, "wrote past the end of a result buffer"
            out.append(r[:n].view(np.uint64))
        return out


def mixed_scenario(dp):
    sc = Scenario(dp)
    topo = sc.build_chain(8, 2)
    spec = config3_spec(n_flows=20000, n_ports=topo["n_ports"] + 4, n_pfx4=3000, n_pfx6=1000, lpm_permille=700,
                        n_acl=300, hit_permille=600)
    install_flows(sc, spec)
    install_lpm_acl(sc, spec)
    dp.commit()
    return sc, spec


@pytest.mark.parametrize("flags", [0, consts.F_COUNTERS])
def test_ragged_scattered_batches(torch_cuda, flags):
    dp = sfc.Dataplane(0)
    sc, spec = mixed_scenario(dp)
    pool = Pool(torch_cuda, spec)
    for n, skew in [(4096, 0), (1, 3), (0, 0), (255, 1), (256, 0), (257, 7), (31, 2), (5000, 5), (100000, 0), (513, 1), (4096, 9)]:
        pool.add(n, skew)
    order = [3, 0, 10, 8, 1, 2, 5, 4, 7, 6, 9]  # not in address order
    dp.classify_batches([pool.batches[i] for i in order], flags, torch_cuda.cuda.current_stream().cuda_stream)
    got = pool.results()
    total = np.zeros(consts.CNT_TOTAL, dtype=np.uint64)
    for g, h, p in zip(got, pool.hdr_np, pool.port_np):
        if not p.size:
            continue
        want, cnt = sc.orc.classify(h, p, counters=True)
        assert_same(g, want, h)
        total += cnt
    if flags & consts.F_COUNTERS:
        assert np.array_equal(dp.counters_read(), total)


def test_config3_batches_in_one_launch(torch_cuda):
    """configs[2] as it is worded: 1M-flow table, 4096-packet batches — 256 of them per call."""
    t = torch_cuda
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    topo = sc.build_chain(8, 4)
    spec = config3_spec(n_flows=1 << 20, n_ports=topo["n_ports"])
    install_flows(sc, spec)
    dp.commit()
    nb, b = 256, 4096
    hdr = t.empty((nb * b, 64), dtype=t.uint8, device="cuda")
    ports = t.empty(nb * b, dtype=t.int16, device="cuda")
    st = t.cuda.current_stream().cuda_stream
    dp.gen_packets(spec, 0, nb * b, hdr.data_ptr(), ports.data_ptr(), st)
    one = t.zeros(nb * b, dtype=t.int64, device="cuda")
    dp.classify(hdr.data_ptr(), ports.data_ptr(), one.data_ptr(), nb * b, 0, st)
    many = t.zeros(nb * b, dtype=t.int64, device="cuda")
    perm = np.random.default_rng(3).permutation(nb)
    batches = [(hdr.data_ptr() + int(i) * b * 64, ports.data_ptr() + int(i) * b * 2, many.data_ptr() + int(i) * b * 8, b) for i in perm]
    for flags in (0, consts.F_OVERLAP, consts.F_OVERLAP):
        many.zero_()
        dp.classify_batches(batches, flags, st)
        t.cuda.synchronize()
        assert t.equal(one, many)
    want = sc.orc.classify(hdr[: 64 * b].cpu().numpy(), ports[: 64 * b].cpu().numpy().view(np.uint16), threads=8)
    assert_same(many[: 64 * b].cpu().numpy().view(np.uint64), want)


def test_lists_longer_than_one_launch(torch_cuda):
    t = torch_cuda
    dp = sfc.Dataplane(0)
    sc, spec = mixed_scenario(dp)
    nb = 1300  # > 512 descriptors: three launches
    sizes = np.random.default_rng(5).integers(1, 200, nb)
    offs = np.concatenate([[0], np.cumsum(sizes)])
    n = int(offs[-1])
    hdr_h, port_h = sfc.gen_packets_host(spec, 7, n)
    hdr = t.from_numpy(np.array(hdr_h, dtype=np.uint8)).cuda()
    ports = t.from_numpy(np.array(port_h, dtype=np.uint16).view(np.int16)).cuda()
    res = t.zeros(n, dtype=t.int64, device="cuda")
    batches = [(hdr.data_ptr() + int(offs[i]) * 64, ports.data_ptr() + int(offs[i]) * 2, res.data_ptr() + int(offs[i]) * 8, int(sizes[i]))
               for i in range(nb)]
    before = dp.launch_count
    dp.classify_batches(batches, 0, t.cuda.current_stream().cuda_stream)
    t.cuda.synchronize()
    assert dp.launch_count - before == 3
    assert_same(res.cpu().numpy().view(np.uint64), sc.orc.classify(hdr_h, port_h), hdr_h)


def test_bad_descriptors_are_rejected_before_any_work(torch_cuda):
    t = torch_cuda
    dp = sfc.Dataplane(0)
    sc, spec = mixed_scenario(dp)
    pool = Pool(t, spec)
    pool.add(1000)
    pool.add(1000)
    good = pool.batches[0]
    hp, pp, rp, n = pool.batches[1]
    st = t.cuda.current_stream().cuda_stream
    for bad in [(hp + 16, pp, rp, n), (0, pp, rp, n), (hp, 0, rp, n), (hp, pp, 0, n)]:
        with pytest.raises(sfc.SfcError) as e:
            dp.classify_batches([good, bad], 0, st)
        assert e.value.code == -errno.EINVAL
    t.cuda.synchronize()
    assert all((r.view(np.int64) == -1).all() for r in pool.results()), "a rejected call must not have launched"
    dp.classify_batches([], 0, st)  # empty list: nothing to do
    dp.classify_batches([(hp, pp, rp, 0)], 0, st)  # only an empty batch
