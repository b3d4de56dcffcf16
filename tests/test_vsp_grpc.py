"""The gRPC front door: a client doing what the daemon's GrpcPlugin does
(internal/daemon/plugin/vendorplugin.go:183-265) against the B200 VSP on a unix socket — the object
sequence of the reference's hostsidemanager_test.go / dpusidemanager flows — must leave the rule set
of tests/golden/replay_rules.json in the dataplane."""
import importlib
import json
import os

import grpc
import pytest

from common import ROOT, sfc

vs = importlib.import_module("dpu-operator_b200.vsp_server")
G = json.load(open(os.path.join(ROOT, "tests", "golden", "replay_rules.json")))
mac = lambda s: bytes(int(x, 16) for x in s.split(":"))  # noqa: E731


@pytest.fixture()
def vsp(tmp_path):
    addr = f"unix:{tmp_path}/vendor-plugin.sock"
    server, impl = vs.serve(addr, device=-1, rpm_port=G["rpm_port"])  # offline dataplane: rule staging only
    client = vs.VspClient(addr)
    yield client, impl
    client.close()
    server.stop(0)


def bridge_port_request(pf, vf, mac_str):
    # HostSideManager.CreateBridgePort (internal/daemon/hostsidemanager.go:48-74)
    req = vs.msg(f"{vs.OPI}.CreateBridgePortRequest")()
    req.bridge_port.name = f"host{pf}-{vf}"
    req.bridge_port.spec.ptype = 1  # ACCESS
    req.bridge_port.spec.mac_address = mac(mac_str)
    req.bridge_port.spec.logical_bridges.append(str(vf + 2))
    return req


def test_daemon_call_sequence_programs_the_26_rules(vsp):
    client, impl = vsp
    ipport = client.Init(vs.msg("Vendor.InitRequest")(dpu_mode=True, dpu_identifier="b200-0"))
    assert (ipport.ip, ipport.port) == ("fe80::1", 8085)
    devs = client.GetDevices(vs.msg("Vendor.Empty")())
    assert sorted(devs.devices) == ["nf_interface0", "nf_interface1"] and devs.devices["nf_interface0"].health == "Healthy"
    for i, m in enumerate(G["vf_macs"]):  # one CNI ADD on the host => exactly one CreateBridgePort (hostsidemanager_test.go:256-261)
        bp = client.CreateBridgePort(bridge_port_request(0, i, m))
        assert bp.name == f"bridge_port/host0-{i}" and bp.spec.mac_address == mac(m) and list(bp.spec.logical_bridges) == [str(i + 2)]
    client.CreateNetworkFunction(vs.msg("Vendor.NFRequest")(input=G["nf_macs"][0], output=G["nf_macs"][1]))
    assert impl.vsp.flow_dump() == G["rules_ports_then_nf"]
    client.DeleteBridgePort(vs.msg(f"{vs.OPI}.DeleteBridgePortRequest")(name=G["deleted_bridge_port"]))
    assert impl.vsp.flow_dump() == G["rules_after_delete_host0_3"]
    client.DeleteNetworkFunction(vs.msg("Vendor.NFRequest")(input=G["nf_macs"][0], output=G["nf_macs"][1]))
    assert impl.vsp.flow_dump() == []


def test_errors_come_back_as_grpc_status_with_the_reference_text(vsp):
    client, _ = vsp
    with pytest.raises(grpc.RpcError) as e:
        client.GetDevices(vs.msg("Vendor.Empty")())  # before Init: "device Store is empty" (marvell/main.go:625-627)
    assert e.value.code() == grpc.StatusCode.INTERNAL and "device Store is empty" in e.value.details()
    client.Init(vs.msg("Vendor.InitRequest")(dpu_mode=True))
    with pytest.raises(grpc.RpcError) as e:
        client.SetNumVfs(vs.msg("Vendor.VfCount")(vf_cnt=8))
    assert "not supported in DPU Mode" in e.value.details()
    req = bridge_port_request(0, 0, G["vf_macs"][0])
    req.bridge_port.name = "eth0"
    with pytest.raises(grpc.RpcError) as e:
        client.CreateBridgePort(req)
    assert e.value.code() == grpc.StatusCode.INTERNAL and "no VFId Match Found" in e.value.details()


def test_host_mode(vsp):
    client, impl = vsp
    ipport = client.Init(vs.msg("Vendor.InitRequest")(dpu_mode=False))
    assert ipport.ip == "fe80::2"
    assert client.SetNumVfs(vs.msg("Vendor.VfCount")(vf_cnt=8)).vf_cnt == 8  # dpudevicehandler.go:89
    assert len(client.GetDevices(vs.msg("Vendor.Empty")()).devices) == 8


def test_wire_format_matches_the_reference_field_numbers():
    # NFRequest{1 input, 2 output} (dpu-api/api.proto:26-29); BridgePortSpec{1 mac bytes, 2 ptype, 3 logical_bridges}
    assert vs.msg("Vendor.NFRequest")(input="a", output="b").SerializeToString() == b"\n\x01a\x12\x01b"
    spec = vs.msg(f"{vs.OPI}.BridgePortSpec")(mac_address=b"\x00\x11\x22\x33\x44\x55", ptype=1)
    spec.logical_bridges.append("7")
    assert spec.SerializeToString() == b"\n\x06\x00\x11\x22\x33\x44\x55\x10\x01\x1a\x017"
    assert vs.msg("Vendor.IpPort")(ip="x", port=8085).SerializeToString() == b"\n\x01x\x10\x95?"


def test_hand_built_descriptors_equal_the_reference_schemas():
    """Every message, field number, field type, service and RPC of dpu-api/api.proto, and the OPI bridge-port
    messages the daemon sends, as tests/golden/make_proto_fixture.py read them out of the reference — compared with
    the descriptors vsp_server.py declares by hand and with the methods it serves."""
    from google.protobuf import descriptor_pb2
    W = json.load(open(os.path.join(ROOT, "tests", "golden", "wire_schema.json")))
    T = descriptor_pb2.FieldDescriptorProto
    scalar = {"bool": T.TYPE_BOOL, "string": T.TYPE_STRING, "int32": T.TYPE_INT32}
    assert W["package"] == "Vendor"
    for name, fields in W["messages"].items():
        d = vs._POOL.FindMessageTypeByName(f"Vendor.{name}")
        assert sorted(f.name for f in d.fields) == sorted(f["name"] for f in fields), name
        for f in fields:
            fd = d.fields_by_name[f["name"]]
            assert fd.number == f["number"], (name, f)
            if f["type"].startswith("map<"):
                k, v = f["type"][4:-1].split(",")
                assert fd.message_type.GetOptions().map_entry and fd.is_repeated
                assert fd.message_type.fields_by_name["key"].type == scalar[k]
                assert fd.message_type.fields_by_name["value"].message_type.full_name == f"Vendor.{v}"
            elif f["type"] in scalar:
                assert fd.type == scalar[f["type"]], (name, f)
            else:
                assert fd.type == T.TYPE_MESSAGE and fd.message_type.full_name == f"Vendor.{f['type']}", (name, f)
    served = {(s, m): io for (s, m), io in vs.METHODS.items()}
    for service, rpcs in W["services"].items():
        for rpc, req, resp in rpcs:
            assert served.pop((f"Vendor.{service}", rpc)) == (f"Vendor.{req}", f"Vendor.{resp}")
    assert W["opi_package"] == vs.OPI
    wire = {"bytes": (T.TYPE_BYTES, T.TYPE_STRING, T.TYPE_MESSAGE), "varint": (T.TYPE_BOOL, T.TYPE_ENUM, T.TYPE_INT32)}
    for name, fields in W["opi_messages"].items():
        d = vs._POOL.FindMessageTypeByName(f"{vs.OPI}.{name}")
        for f in fields:
            if f["name"] == "components":  # BridgePortStatus.components: never set by the reference VSPs, not declared here
                continue
            fd = d.fields_by_name[f["name"]]
            assert fd.number == f["number"] and fd.type in wire[f["wire"]], (name, f)
            assert fd.is_repeated == (f["label"] == "rep"), (name, f)
            if f["enum"]:
                assert fd.enum_type.full_name == f["enum"]
    # what is left are the two OPI methods the daemon calls; the reference registers the whole BridgePortService
    assert sorted(m for (s, m) in served) == ["CreateBridgePort", "DeleteBridgePort"]
    assert {"CreateBridgePort", "DeleteBridgePort"} <= set(W["opi_bridge_port_methods"])


def test_stat_surface_and_health(vsp):
    """GetDevices reports whether the dataplane answers, and B200Vsp.stats() reads its hit counters back."""
    client, server = vsp
    assert server.stats() is None  # no dataplane before Init
    client.Init(vs.msg("Vendor.InitRequest")(dpu_mode=True))
    st = server.stats()
    assert st["forward"] == 0 and set(st["by_stage"]) == {"miss", "acl", "flow", "l2_exact", "l2_wildcard", "lpm4", "lpm6"}
    devs = client.GetDevices(vs.msg("Vendor.Empty")())
    assert {d.health for d in devs.devices.values()} == {"Healthy"}


def test_daemon_side_mac_pairing_drives_the_nf_rpcs(vsp):
    """dpusidemanager.go:104-139: two CNI ADDs in one netns make ONE CreateNetworkFunction(first MAC, second MAC); the
    first DEL makes one DeleteNetworkFunction with the same pair, the second DEL none."""
    client, impl = vsp
    client.Init(vs.msg("Vendor.InitRequest")(dpu_mode=True))
    calls = []

    class Spy:
        def CreateNetworkFunction(self, a, b):
            calls.append(("create", a, b))
            client.CreateNetworkFunction(vs.msg("Vendor.NFRequest")(input=a, output=b))

        def DeleteNetworkFunction(self, a, b):
            calls.append(("delete", a, b))
            client.DeleteNetworkFunction(vs.msg("Vendor.NFRequest")(input=a, output=b))

    pair = vs.NfMacPairer(Spy())
    m0, m1 = G["nf_macs"]
    pair.cni_add("/var/run/netns/nf", m0)
    assert calls == []
    pair.cni_add("/var/run/netns/other", "02:00:00:00:09:09")  # another pod's first interface does not complete the pair
    pair.cni_add("/var/run/netns/nf", m1)
    assert calls == [("create", m0, m1)]
    assert any(",actions=output:" + G["rpm_port"] in r for r in impl.vsp.flow_dump())  # the NF's wire rules are programmed
    pair.cni_del("/var/run/netns/nf")
    pair.cni_del("/var/run/netns/nf")
    assert calls == [("create", m0, m1), ("delete", m0, m1)] and pair.mac_store["/var/run/netns/nf"] == []
