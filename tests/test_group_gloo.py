"""Multi-rank host logic on CPU (gloo, world_size 2): packet sharding, table-image broadcast and
counter reduction of dpu-operator_b200/group.py, with offline handles standing in for the GPUs'."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import importlib

    import torch.distributed as dist

    from common import Scenario, config3_spec, install_flows, sfc

    group = importlib.import_module("dpu-operator_b200.group")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dp = sfc.Dataplane(-1)
    spec = config3_spec(n_flows=5000, n_ports=13)
    if rank == 0:  # only rank 0 holds the rules; the others receive the image
        sc = Scenario(dp)
        sc.build_chain(8, 2)
        install_flows(sc, spec)
        dp.commit()
    group.broadcast_table(dp, rank, world)
    image = dp.table_export()
    # every rank classifies its contiguous share (here: with the oracle, as the GPUs are absent)
    n = 10007
    lo, hi = group.shard_range(n, rank, world)
    sc2 = Scenario(sfc.Dataplane(-1))
    sc2.build_chain(8, 2)
    install_flows(sc2, spec)
    hdr, ports = sfc.gen_packets_host(spec, lo, hi - lo)
    res, cnt = sc2.orc.classify(hdr, ports, counters=True)
    total = group.reduce_counters_host(cnt)
    np.save(os.path.join(out_dir, f"image{rank}.npy"), image)
    np.save(os.path.join(out_dir, f"res{rank}.npy"), res)
    np.save(os.path.join(out_dir, f"total{rank}.npy"), total)
    dist.destroy_process_group()


def test_two_ranks_broadcast_shard_and_reduce(tmp_path):
    import torch.multiprocessing as mp

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from common import Scenario, config3_spec, consts, install_flows, sfc

    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    img0, img1 = np.load(tmp_path / "image0.npy"), np.load(tmp_path / "image1.npy")
    assert img0.size > 512 and np.array_equal(img0, img1), "rank 1 did not receive rank 0's table image"
    # the shards, concatenated, equal one rank doing everything; the reduced counters equal its counters
    spec = config3_spec(n_flows=5000, n_ports=13)
    sc = Scenario(sfc.Dataplane(-1))
    sc.build_chain(8, 2)
    install_flows(sc, spec)
    hdr, ports = sfc.gen_packets_host(spec, 0, 10007)
    want, cnt = sc.orc.classify(hdr, ports, counters=True)
    got = np.concatenate([np.load(tmp_path / "res0.npy"), np.load(tmp_path / "res1.npy")])
    assert np.array_equal(got, want)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"total{r}.npy"), cnt)
    assert cnt[consts.CNT_ACTION0:consts.CNT_ACTION0 + 3].sum() == 10007


@pytest.mark.parametrize("n,world", [(10, 3), (0, 4), (1 << 28, 8), (7, 8)])
def test_shard_range_partitions_exactly(n, world):
    import importlib

    sys.path.insert(0, ROOT)
    group = importlib.import_module("dpu-operator_b200.group")
    spans = [group.shard_range(n, r, world) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
    assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= 1
