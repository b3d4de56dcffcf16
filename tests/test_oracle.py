"""CPU tests of the oracle (no GPU): the C++ oracle against the independent Python restatement,
against the committed golden fixture, and against properties of the semantics."""
import json
import os

import numpy as np

from common import ROOT, Oracle, Scenario, config3_spec, consts, install_flows, install_lpm_acl, sfc
from packets import handmade_cases

GOLDEN = os.path.join(ROOT, "tests", "golden", "handmade_cases.json")


def test_cpp_oracle_equals_python_oracle_on_handmade_cases():
    sc = Scenario(sfc.Dataplane(-1), with_py=True)
    hdr, ports = handmade_cases(sc)
    a = sc.orc.classify(hdr, ports)
    b = np.array([sc.py.classify(bytes(hdr[i]), int(ports[i])) for i in range(len(ports))], dtype=np.uint64)
    assert np.array_equal(a, b), np.nonzero(a != b)[0]
    assert set(sfc.result_fields(a)[3].tolist()) == set(range(7))  # every stage decided something


def test_golden_fixture():
    """tests/golden/handmade_cases.json was written by tests/golden/make_golden.py from the Python
    oracle and then reviewed by hand (each entry carries a note saying why the answer is right)."""
    g = json.load(open(GOLDEN))
    sc = Scenario(sfc.Dataplane(-1))
    hdr, ports = handmade_cases(sc)
    assert len(g["cases"]) == len(ports)
    got = sc.orc.classify(hdr, ports)
    for i, c in enumerate(g["cases"]):
        assert bytes(hdr[i]).hex() == c["hdr"] and int(ports[i]) == c["in_port"], f"fixture input {i} drifted"
        assert int(got[i]) == int(c["result"], 16), f"case {i} ({c['note']}): got {int(got[i]):#018x} want {c['result']}"


def test_cpp_oracle_equals_python_oracle_on_seeded_traffic():
    sc = Scenario(sfc.Dataplane(-1), with_py=True)
    topo = sc.build_chain(4, 2)
    spec = config3_spec(n_flows=3000, n_ports=topo["n_ports"] + 3, n_pfx4=400, n_pfx6=150, lpm_permille=700, n_acl=60,
                        hit_permille=500)
    install_flows(sc, spec)
    install_lpm_acl(sc, spec)
    hdr, ports = sfc.gen_packets_host(spec, 0, 4000)
    a = sc.orc.classify(hdr, ports)
    b = np.array([sc.py.classify(bytes(hdr[i]), int(ports[i])) for i in range(len(ports))], dtype=np.uint64)
    assert np.array_equal(a, b), np.nonzero(a != b)[0][:10]
    assert len(set(sfc.result_fields(a)[3].tolist())) == 7


def test_threads_do_not_change_results():
    sc = Scenario(sfc.Dataplane(-1))
    spec = config3_spec(n_flows=2000, n_ports=8)
    install_flows(sc, spec)
    hdr, ports = sfc.gen_packets_host(spec, 0, 10001)
    assert np.array_equal(sc.orc.classify(hdr, ports, threads=1), sc.orc.classify(hdr, ports, threads=5))


def test_reachability_matrix():
    """The reference's e2e suite (e2e_test/e2e_test.go:406-513) pings pod<->pod, pod<->NF and
    NF/pod<->external with the chain installed.  Restated per hop: with the Marvell steering set
    (marvell/main.go:382-403, 519-560) every one of those frames must be FORWARDed to the next
    port on its path."""
    sc = Scenario(sfc.Dataplane(-1))
    t = sc.build_chain(n_vfs=2, n_nfs=1)
    vf0, vf1 = t["vfs"]
    inp, out, wire = t["nf_in"][0], t["nf_out"][0], t["wire"]
    m0, m1, ext = sfc.gen_mac(0), sfc.gen_mac(1), bytes.fromhex("0200000000fe")
    from packets import SMAC, eth

    arp = bytes(28)
    hops = [  # (dst mac, in_port, expected out_port, expected stage)
        (m1, vf0, inp, consts.STAGE_L2_WILD),    # pod1 -> pod2 enters the NF
        (m1, out, out, consts.STAGE_L2_EXACT),   # NF egress hairpins traffic for a local VF (in_port action)
        (m1, inp, vf1, consts.STAGE_L2_EXACT),   # ... and the NF ingress side delivers it to pod2
        (m0, inp, vf0, consts.STAGE_L2_EXACT),   # reply reaches pod1
        (ext, vf0, inp, consts.STAGE_L2_WILD),   # pod -> external goes through the NF
        (ext, out, wire, consts.STAGE_L2_WILD),  # NF -> wire
        (m0, wire, out, consts.STAGE_L2_WILD),   # external -> NF
    ]
    hdr = np.frombuffer(b"".join(eth(d, SMAC, 0x0806, arp) for d, *_ in hops), np.uint8).reshape(-1, 64)
    ports = np.array([h[1] for h in hops], np.uint16)
    act, nf, port, stage, _ = sfc.result_fields(sc.orc.classify(hdr, ports))
    assert act.tolist() == [consts.ACT_FORWARD] * len(hops)
    assert port.tolist() == [h[2] for h in hops]
    assert stage.tolist() == [h[3] for h in hops]
    assert nf[0] == 0 and nf[5] == consts.NF_NONE  # NF ports belong to NF 0, the wire to none


def test_l2_delete_is_loose_like_ovs_del_flows():
    # `ovs-ofctl del-flows br in_port=S` removes MAC-qualified flows too (ovs-dp/ovsdp.go:143-162)
    sc = Scenario(sfc.Dataplane(-1))
    t = sc.build_chain(2, 1)
    inp = t["nf_in"][0]
    from packets import SMAC, eth

    frame = np.frombuffer(eth(sfc.gen_mac(0), SMAC, 0x0806, bytes(28)), np.uint8).reshape(1, 64)
    p = np.array([inp], np.uint16)
    assert sfc.result_fields(sc.orc.classify(frame, p))[3][0] == consts.STAGE_L2_EXACT
    sc.l2_del(inp, None)
    assert sc.orc.classify(frame, p)[0] == consts.RESULT_MISS


def test_the_two_writings_of_the_workload_agree_byte_for_byte():
    """oracle/workload.cpp and the product's generator (csrc/gen.h) implement one specification twice; flows,
    packets, prefixes, MACs and ACL rules must come out identical for every traffic mix the configs use."""
    from common import OracleGen, config2_spec, config3_spec, config4_spec

    specs = [config2_spec(), config3_spec(n_flows=50000, n_ports=17), config3_spec(n_flows=(1 << 20) + 12345, n_ports=9, zipf=1),
             config4_spec(n_pfx4=3000, n_pfx6=900, n_acl=300), config3_spec(n_flows=777, v6_permille=1000, vlan_permille=1000, zipf=0)]
    for spec in specs:
        nf = min(int(spec["n_flows"][0]), 20000)
        if nf:
            k1, a1 = sfc.gen_flows(spec, 3, nf - 3)
            k2, a2 = OracleGen.flows(spec, 3, nf - 3)
            assert k1.tobytes() == k2.tobytes() and a1.tobytes() == a2.tobytes()
        h1, p1 = sfc.gen_packets_host(spec, 1000, 30000)
        h2, p2 = OracleGen.packets(spec, 1000, 30000)
        assert np.array_equal(p1, p2)
        assert np.array_equal(h1, h2), f"first differing packet: {int(np.nonzero((h1 != h2).any(axis=1))[0][0])}"
        for af, n in ((4, int(spec["n_pfx4"][0])), (6, int(spec["n_pfx6"][0]))):
            mine = OracleGen.prefixes(spec, af, min(n, 500))
            theirs = []
            for j in range(min(n, 500)):
                import ctypes as C
                buf, ln, port = C.create_string_buffer(16), C.c_int(), C.c_uint16()
                sfc.capi.lib.sfc_gen_prefix(spec.ctypes.data, af, j, C.cast(buf, C.c_void_p), C.byref(ln), C.byref(port))
                theirs.append((buf.raw[: 4 if af == 4 else 16], ln.value, port.value))
            assert mine == theirs
        na = int(spec["n_acl"][0])
        if na:
            assert OracleGen.acl_rules(spec, na).tobytes() == sfc.gen_acl_rules(spec, na).tobytes()
    assert all(OracleGen.mac(i) == sfc.gen_mac(i) for i in (0, 1, 7, 255, 256, 70000))


def test_reference_arm_of_the_bench_loads_no_product_code():
    """`bench.py --impl reference` times the CPU oracle on inputs from the oracle's own generator.  With the product
    library made unloadable the arm must still run (importing the package would raise), and its line must carry the
    keys the driver reads."""
    import subprocess
    import sys

    env = dict(os.environ, B200SFC_LIB="/nonexistent/libb200sfc.so")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "Mpps" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0
    assert line["config"]["workload"].startswith("configs[1]")
    # and the GPU arm names the same workload (the driver compares the two strings)
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert src.count('"workload": WORKLOAD') == 2


def test_bench_stager_numbers_rules_like_the_library():
    """bench.py's ranks other than 0 stage rules into their oracle without a library handle and must hand out the ids
    the library hands out on rank 0 (ports, L2 rules, prefixes, ACL rules, flows): run the stager against an offline
    handle — its asserts compare every id — and check that the handle-less twin ends with the same counts."""
    import importlib.util
    import sys

    spec_ = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec_)
    spec_.loader.exec_module(bench)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc

    with_dp = bench.Stager(sfc.Dataplane(-1), orc, sfc)
    without = bench.Stager(None, orc, sfc)
    small = dict(bench.CFG3_SPEC, n_pfx4=600, n_pfx6=200, n_acl=50)
    for st in (with_dp, without):
        n_ports = st.chain(8, 4)
        st.flows(sfc.make_gen_spec(**bench.cfg2_spec_kw(5000, n_ports)), 5000, chunk=2048)
        st.lpm_acl(sfc.make_gen_spec(**small))
    with_dp.dp.commit()
    assert (with_dp.n_ports, with_dp.n_l2, with_dp.n_lpm, with_dp.n_acl, with_dp.n_flows) == \
           (without.n_ports, without.n_l2, without.n_lpm, without.n_acl, without.n_flows) == (17, 8 * 6 + 5, with_dp.n_lpm, 50, 5000)
    # and the two oracles answer alike
    pk = sfc.make_gen_spec(**bench.cfg2_spec_kw(5000, 17))
    hdr, ports = orc.OracleGen.packets(pk, 0, 20000)
    assert np.array_equal(with_dp.orc.classify(hdr, ports), without.orc.classify(hdr, ports))
