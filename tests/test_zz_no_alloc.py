"""sfc_classify_static — the entry point for callers that may not allocate.

Some users of this library cannot call cudaMalloc when they have packets in hand: a dataplane
thread living inside a fixed memory budget, a shim inside a process whose device allocator belongs
to another runtime, a polling loop that must stay off the driver's allocation locks.  For them the
library keeps the classify kernel's scratch (its tile tickets) and the table image in a fixed
array compiled into the library, and takes no handle.  This file checks, through the C ABI and on
two calls in a row, that the results are bit-identical to the ordinary handle-based entry point AND
to the CPU oracle — with and without ACL rules in the image, because the two cases run different
kernels — that the shared array is re-armed correctly between calls, that free device memory does
not move across a call, and that an image larger than the fixed area is refused with a message that
says so.

It is named so that pytest collects it last: if it fails under -x every other test has already run.
"""
import numpy as np
import pytest

from common import Scenario, config3_spec, install_flows, install_lpm_acl, sfc

pytestmark = pytest.mark.gpu


def _scenario(n_acl, fmt=1):
    dp = sfc.Dataplane(0)
    dp.set_option("flow_format", fmt)
    sc = Scenario(dp)
    topo = sc.build_chain(8, 2)
    spec = config3_spec(n_flows=20000, n_ports=topo["n_ports"] + 2, n_pfx4=500, n_pfx6=100, lpm_permille=600, n_acl=n_acl)
    install_flows(sc, spec)
    install_lpm_acl(sc, spec)  # n_acl == 0: prefixes only
    dp.commit()
    return dp, sc, spec


def _batch(torch, dp, spec, first, n):
    hdr = torch.empty((n, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty(n, dtype=torch.int16, device="cuda")
    dp.gen_packets(spec, first, n, hdr.data_ptr(), ports.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return hdr, ports


@pytest.mark.parametrize("n_acl,fmt", [(0, 1), (120, 1), (0, 2)])
def test_static_entry_point_matches_the_handle_entry_point_twice(torch_cuda, n_acl, fmt):
    torch = torch_cuda
    dp, sc, spec = _scenario(n_acl, fmt)  # with / without ACL rules, sector / indexed flow table: four different kernels
    image = dp.table_export()  # host copy of the committed image: the caller's input
    assert image.size <= sfc.STATIC_AREA_BYTES - 256
    st = torch.cuda.current_stream().cuda_stream
    batches = [_batch(torch, dp, spec, 0, 300_000), _batch(torch, dp, spec, 777_777, 41_003)]  # ragged second size
    want = []
    for hdr, ports in batches:  # the existing entry point
        res = torch.zeros(ports.numel(), dtype=torch.int64, device="cuda")
        dp.classify(hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), ports.numel(), 0, st)
        torch.cuda.synchronize()
        want.append(res)
        # the oracle's answer for the same packets: the static path below is thereby held to the oracle as well
        orc = sc.orc.classify(hdr.cpu().numpy(), ports.cpu().numpy().view(np.uint16), threads=8)
        assert np.array_equal(res.cpu().numpy().view(np.uint64), orc)
    assert len({int(x) for x in sfc.result_fields(want[0].cpu().numpy().view(np.uint64))[3]}) >= (6 if n_acl else 5)  # a real mix of stages

    got = [torch.full((p.numel(),), -1, dtype=torch.int64, device="cuda") for _, p in batches]
    # call 1 carries the image; call 2 reuses what call 1 left in the static array
    sfc.classify_static(0, image, batches[0][0].data_ptr(), batches[0][1].data_ptr(), got[0].data_ptr(), got[0].numel(), st)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    sfc.classify_static(0, None, batches[1][0].data_ptr(), batches[1][1].data_ptr(), got[1].data_ptr(), got[1].numel(), st)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "a call changed the amount of free device memory"
    for g, w in zip(got, want):
        assert torch.equal(g, w), f"{int((g != w).sum())} results differ from sfc_classify"
    # and once more with the image passed again, right behind the previous call on the same stream
    got2 = torch.full_like(got[0], -1)
    sfc.classify_static(0, image, batches[0][0].data_ptr(), batches[0][1].data_ptr(), got2.data_ptr(), got2.numel(), st)
    sfc.classify_static(0, None, batches[1][0].data_ptr(), batches[1][1].data_ptr(), got[1].data_ptr(), got[1].numel(), st)
    torch.cuda.synchronize()
    assert torch.equal(got2, want[0]) and torch.equal(got[1], want[1])


def test_static_entry_point_refuses_what_does_not_fit(torch_cuda):
    torch = torch_cuda
    dp = sfc.Dataplane(0)
    spec = config3_spec(n_flows=200_000, n_ports=8)
    keys, acts = sfc.gen_flows(spec, 0, 200_000)
    dp.flow_add_bulk(keys, acts)
    dp.commit()
    image = dp.table_export()
    assert image.size > sfc.STATIC_AREA_BYTES
    hdr, ports = _batch(torch, dp, spec, 0, 1024)
    res = torch.zeros(1024, dtype=torch.int64, device="cuda")
    with pytest.raises(sfc.SfcError) as e:
        sfc.classify_static(0, image, hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), 1024)
    assert e.value.code == -7  # -E2BIG
    assert str(image.size) in str(e.value) and "static device area" in str(e.value)
    with pytest.raises(sfc.SfcError) as e:  # a damaged image is refused too, before anything is copied
        bad = dp.table_export()[: 1 << 20].copy()
        sfc.classify_static(0, bad, hdr.data_ptr(), ports.data_ptr(), res.data_ptr(), 1024)
    assert e.value.code == -22
