"""sfc_group_* — several GPUs driven from one process through the C ABI: rank 0's table
replicated with ncclBroadcast, packets sharded, counters summed with ncclAllReduce.  The multi-GPU test needs two
GPUs and skips otherwise (run it with `gpurun --gpus 2`); the single-GPU group and the error paths run anywhere."""
import errno

import numpy as np
import pytest

from common import Scenario, config3_spec, consts, install_flows, install_lpm_acl, sfc


def test_no_gpu_means_no_group():
    import ctypes
    g = ctypes.c_void_p()
    rc = sfc.lib.sfc_group_create(0, None, ctypes.byref(g))
    assert rc == -errno.EINVAL and not g.value
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # noqa: BLE001
        has_gpu = False
    if not has_gpu:
        assert sfc.lib.sfc_group_create(2, None, ctypes.byref(g)) == -errno.ENODEV and not g.value
    assert sfc.lib.sfc_group_size(None) == -errno.EINVAL and sfc.lib.sfc_group_handle(None, 0) is None


def _scenario(dp):
    sc = Scenario(dp)
    topo = sc.build_chain(8, 2)
    spec = config3_spec(n_flows=50000, n_ports=topo["n_ports"] + 2, n_pfx4=2000, n_pfx6=500, lpm_permille=600, n_acl=100, hit_permille=700)
    install_flows(sc, spec)
    install_lpm_acl(sc, spec)
    dp.commit()
    return sc, spec


@pytest.mark.gpu
def test_group_of_one(torch_cuda):
    g = sfc.Group(1)
    sc, spec = _scenario(g[0])
    assert g.broadcast_table() == 0.0 and g.reduce_counters() == 0.0  # nothing to do, no NCCL needed
    hdr, ports = sfc.gen_packets_host(spec, 0, 10000)
    assert np.array_equal(g[0].classify_host(hdr, ports), sc.orc.classify(hdr, ports))
    g.close()


@pytest.mark.gpu
def test_broadcast_shard_reduce_across_gpus(torch_cuda):
    t = torch_cuda
    n_gpus = t.cuda.device_count()
    if n_gpus < 2:
        pytest.skip("needs two GPUs")
    n_gpus = min(n_gpus, 4)
    with pytest.raises(sfc.SfcError):
        sfc.Group(2, devices=[0, 0])  # one handle per GPU
    g = sfc.Group(n_gpus)
    sc, spec = _scenario(g[0])
    with_table = g[0].table_size
    assert all(g[r].table_size == 0 for r in range(1, n_gpus))
    ms = g.broadcast_table()
    assert ms > 0 and all(g[r].table_size == with_table and g[r].table_epoch == g[0].table_epoch for r in range(n_gpus))
    n = 1 << 18
    hdr, ports = sfc.gen_packets_host(spec, 0, n)
    want, cnt = sc.orc.classify(hdr, ports, counters=True)
    shard = n // n_gpus
    outs = []
    for r in range(n_gpus):
        lo, hi = r * shard, (n if r == n_gpus - 1 else (r + 1) * shard)
        with t.cuda.device(r):
            d_hdr = t.from_numpy(np.array(hdr[lo:hi])).cuda()
            d_port = t.from_numpy(np.array(ports[lo:hi]).view(np.int16)).cuda()
            res = t.zeros(hi - lo, dtype=t.int64, device="cuda")
            g[r].classify(d_hdr.data_ptr(), d_port.data_ptr(), res.data_ptr(), hi - lo, consts.F_COUNTERS, t.cuda.current_stream().cuda_stream)
            outs.append((res, d_hdr, d_port))
    got = np.concatenate([o[0].cpu().numpy().view(np.uint64) for o in outs])
    assert np.array_equal(got, want)
    assert g.reduce_counters() > 0
    for r in range(n_gpus):
        assert np.array_equal(g[r].counters_read(), cnt), f"rank {r} does not hold the total"
    # a second image (one more rule) replaces the first everywhere
    p = g[0].port_add("late")
    g[0].l2_rule_add(p, p)
    g[0].commit()
    g.broadcast_table()
    assert all(g[r].table_epoch == g[0].table_epoch for r in range(n_gpus))
    g.close()


@pytest.mark.gpu
def test_c_host_program_drives_the_gpus(torch_cuda, tmp_path):
    """integration/c/group_demo.c: no Python, no CUDA headers — the C ABI alone, as a cgo host would use it."""
    import os
    import shutil
    import subprocess
    if not shutil.which("gcc"):
        pytest.skip("no gcc")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.join(root, "dpu-operator_b200")
    exe = tmp_path / "group_demo"
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", "-I", os.path.join(root, "include"), os.path.join(root, "integration", "c", "group_demo.c"),
                    "-L", libdir, "-lb200sfc", f"-Wl,-rpath,{libdir}", "-o", str(exe)], check=True)
    n = min(torch_cuda.cuda.device_count(), 2)
    out = subprocess.run([str(exe), str(n)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    # NCCL prints its version banner to stdout on some hosts before the program's own line
    assert any(line.startswith(f"group_demo ok: {n} GPUs") for line in out.stdout.splitlines()), out.stdout
