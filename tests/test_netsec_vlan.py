"""the NetSec VSP's only static rule content is "VF i carries VLAN i + 2"
(/root/reference/internal/daemon/vendor-specific-plugins/intel-netsec/main.go:41, 444-445).  Expressed as two ACL
rules per VF — permit (in_port, vlan), deny (in_port) below it — the classifier must drop exactly the frames that
arrive on a VF port with a foreign or missing tag; the image walker (CPU) and the CUDA path (GPU) agree with the oracle."""
import numpy as np
import pytest

from common import Scenario, consts, sfc
from packets import acl, eth, ipv4, ports

VLAN_OFFSET = 2  # intel-netsec/main.go:41


def netsec(sc, n_vfs):
    vf_ports = [sc.dp.port_add(f"host0-{i}") for i in range(n_vfs)]
    uplink = sc.dp.port_add("uplink")
    for i, p in enumerate(vf_ports):
        sc.acl_add(acl(in_port=p, vlan=i + VLAN_OFFSET, priority=2, deny=False))
        sc.acl_add(acl(in_port=p, priority=1, deny=True))
    sc.lpm_add(4, bytes([10, 0, 0, 0]), 8, uplink)  # something for permitted traffic to hit
    return vf_ports, uplink


def frames(vf_ports, uplink):
    mac = bytes(6)
    out = []
    for i, p in enumerate(vf_ports):
        for vlan in (None, 0, 1, i + VLAN_OFFSET, i + VLAN_OFFSET + 1, 4095):
            tags = [] if vlan is None else [(0x8100, vlan)]
            out.append((eth(mac, mac, 0x0800, ipv4(bytes([10, 1, 1, 1]), bytes([10, 2, 2, 2]), 6, ports(1, 2)), tags=tags), p,
                        vlan == i + VLAN_OFFSET))
    out.append((eth(mac, mac, 0x0800, ipv4(bytes([10, 1, 1, 1]), bytes([10, 2, 2, 2]), 6, ports(1, 2)), tags=[(0x8100, 77)]), uplink, True))
    hdr = np.stack([np.frombuffer(f, np.uint8) for f, _, _ in out])
    return hdr, np.array([p for _, p, _ in out], np.uint16), np.array([ok for _, _, ok in out])


def check(result, allowed):
    action, _, _, stage, _ = sfc.result_fields(result)
    assert ((action == consts.ACT_DROP) == ~allowed).all()
    assert (stage[~allowed] == consts.STAGE_ACL).all() and (stage[allowed] == consts.STAGE_LPM4).all()


def test_vlan_isolation_in_the_oracles():
    sc = Scenario(sfc.Dataplane(-1), with_py=True)
    hdr, in_port, allowed = frames(*netsec(sc, 8))
    want = sc.orc.classify(hdr, in_port)
    check(want, allowed)
    assert np.array_equal(np.array([sc.py.classify(bytes(h), int(p)) for h, p in zip(hdr, in_port)], dtype=np.uint64), want)


@pytest.mark.gpu
def test_vlan_isolation_on_the_gpu(torch_cuda):
    dp = sfc.Dataplane(0)
    sc = Scenario(dp)
    hdr, in_port, allowed = frames(*netsec(sc, 8))
    dp.commit()
    got = dp.classify_host(hdr, in_port)
    assert np.array_equal(got, sc.orc.classify(hdr, in_port))
    check(got, allowed)
