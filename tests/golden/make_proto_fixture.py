"""Regenerate tests/golden/wire_schema.json — the message schemas the VSP gRPC front door must speak, read out of the
reference (run in the build container): dpu-api/api.proto (parsed as text; no protoc in this image) and the
`protobuf:"..."` struct tags of the vendored OPI bridge-port messages
(vendor/github.com/opiproject/opi-api/network/evpn-gw/v1alpha1/gen/go/l2_xpu_infra_mgr.pb.go).
tests/test_vsp_grpc.py compares the descriptors dpu-operator_b200/vsp_server.py builds by hand with this fixture.
    python tests/golden/make_proto_fixture.py
"""
import json
import os
import re

REF = "/root/reference"
proto = open(os.path.join(REF, "dpu-api/api.proto")).read()
package = re.search(r"^package (\w+);", proto, re.M).group(1)
messages = {}
for name, body in re.findall(r"message (\w+) \{(.*?)\}", proto, re.S):
    fields = []
    for line in body.strip().splitlines():
        m = re.match(r"\s*(repeated\s+)?(map<\s*\w+\s*,\s*\w+\s*>|\w+)\s+(\w+)\s*=\s*(\d+);", line)
        if m:
            fields.append(dict(name=m.group(3), number=int(m.group(4)), type=re.sub(r"\s+", "", m.group(2)), repeated=bool(m.group(1))))
    messages[name] = fields
services = {}
for name, body in re.findall(r"service (\w+) \{(.*?)\}", proto, re.S):
    services[name] = [list(r) for r in re.findall(r"rpc (\w+)\((\w+)\) returns \((\w+)\);", body)]

pbgo = open(os.path.join(REF, "vendor/github.com/opiproject/opi-api/network/evpn-gw/v1alpha1/gen/go/l2_xpu_infra_mgr.pb.go")).read()
opi = {}
for name in ("BridgePort", "BridgePortSpec", "BridgePortStatus", "CreateBridgePortRequest", "DeleteBridgePortRequest"):
    body = re.search(r"\ntype " + name + r" struct \{(.*?)\n\}", pbgo, re.S).group(1)
    opi[name] = [dict(wire=w, number=int(n), label=l, name=f, enum=(e or None))
                 for w, n, l, f, e in re.findall(r'protobuf:"(\w+),(\d+),(\w+),name=(\w+)(?:,json=\w+)?,proto3(?:,enum=([\w.]+))?"', body)]
opi_package = re.search(r'enum=([\w.]+)\.BridgePortType', pbgo).group(1)
grpc_go = open(os.path.join(REF, "vendor/github.com/opiproject/opi-api/network/evpn-gw/v1alpha1/gen/go/l2_xpu_infra_mgr_grpc.pb.go")).read()
opi_methods = sorted(set(re.findall(r'"/' + re.escape(opi_package) + r'\.BridgePortService/(\w+)"', grpc_go)))
out = dict(source="dpu-api/api.proto + vendored opi-api l2_xpu_infra_mgr.pb.go struct tags", package=package, messages=messages,
           services=services, opi_package=opi_package, opi_messages=opi, opi_bridge_port_methods=opi_methods)
path = os.path.join(os.path.dirname(__file__), "wire_schema.json")
json.dump(out, open(path, "w"), indent=1)
print(f"wrote {path}: {len(messages)} messages, {sum(len(v) for v in services.values())} rpcs, {len(opi)} OPI messages, methods {opi_methods}")
