"""Regenerate tests/golden/replay_rules.json — BASELINE.json configs[0] without a cluster.

configs[0] is "reconcile examples/sfc.yaml + dpu.yaml/host.yaml on envtest/Kind, emitting the rule
set".  Kind/envtest/Go are not available, so this script (run in the build container, where
/root/reference is mounted) replays the call sequence the reconcilers would make and writes the
rule set the reference's OVS backend would have programmed:

  * the inputs come from the reference's own files: examples/sfc.yaml (the chain), examples/dpu.yaml /
    host.yaml (modes), examples/my-pod.yaml (one openshift.io/dpu per workload pod), the VF count
    of internal/daemon/device-handler/dpu-device-handler/dpudevicehandler.go:89 (SetNumVfs(8));
  * the `ovs-ofctl add-flow` format strings are READ OUT OF
    internal/daemon/vendor-specific-plugins/marvell/ovs-dp/ovsdp.go with a regex, not retyped;
  * the order of calls follows marvell/main.go CreateBridgePort (:347-421) and AddNetworkFunction
    (:501-563), transliterated below in a dozen lines of Python.

The fixture is what tests/test_replay.py holds the C++ mirror (csrc/host/) to.  Run:
    python tests/golden/make_replay.py
"""
import json
import os
import re

import yaml

REF = "/root/reference"
OVSDP = os.path.join(REF, "internal/daemon/vendor-specific-plugins/marvell/ovs-dp/ovsdp.go")
HANDLER = os.path.join(REF, "internal/daemon/device-handler/dpu-device-handler/dpudevicehandler.go")

src = open(OVSDP).read()
fmts = re.findall(r'"add-flow", bridgeName, fmt\.Sprintf\("([^"]+)"', src)
assert len(fmts) == 3, fmts
HAIRPIN, MACRULE, WILDCARD = fmts  # order of appearance in AddFlowRuleToDataPlane (:120, :127, :135)
assert HAIRPIN.startswith("priority=100") and WILDCARD.startswith("priority=10,") and MACRULE.startswith("in_port=")
num_vfs = int(re.search(r"SetNumVfs\((\d+)\)", open(HANDLER).read()).group(1))  # dpudevicehandler.go:89

sfc = yaml.safe_load(open(os.path.join(REF, "examples/sfc.yaml")))
dpu = yaml.safe_load(open(os.path.join(REF, "examples/dpu.yaml")))
host = yaml.safe_load(open(os.path.join(REF, "examples/host.yaml")))
pod = yaml.safe_load(open(os.path.join(REF, "examples/my-pod.yaml")))
assert dpu["spec"]["mode"] == "dpu" and host["spec"]["mode"] == "host"
assert pod["spec"]["containers"][0]["resources"]["requests"]["openshift.io/dpu"] == "1"
nfs = [n["name"] for n in sfc["spec"]["networkFunctions"]]

RPM = "rpm0"
vf_macs = [f"00:{i:02x}:00:00:03:14" for i in range(num_vfs - 1)] + ["00:11:22:33:44:55"]  # last: hostsidemanager_test.go fixture MAC
nf_macs = ["02:00:00:00:01:00", "02:00:00:00:01:01"]  # stand-ins for the kernel-assigned veth MACs
dp_ifs = ["dp_interface0", "dp_interface1"]              # main.go:109


def vf_name(pf, vf):  # main.go:321,333
    return f"vf{pf}-{vf + 1}"


def add(rules, in_port, out_port, mac):  # ovsdp.go:113-142
    if mac:
        line = HAIRPIN % (in_port, mac) if in_port == out_port else MACRULE % (in_port, mac, out_port)
    else:
        line = WILDCARD % (in_port, out_port)
    rules.append(line)


def vf_rules(rules, vf, mac, inp, out):  # main.go:383-403 == :519-536
    add(rules, vf, inp, "")
    add(rules, inp, vf, mac)
    add(rules, out, out, mac)


inp, out = dp_ifs
# order A: host pods first (8 bridge ports), then the chain
a = []
for i, m in enumerate(vf_macs):
    vf_rules(a, vf_name(0, i), m, inp, out)
add(a, out, RPM, "")
add(a, RPM, out, "")
# order B: chain first, then the bridge ports (each CreateBridgePort adds its three rules)
b = []
add(b, out, RPM, "")
add(b, RPM, out, "")
for i, m in enumerate(vf_macs):
    vf_rules(b, vf_name(0, i), m, inp, out)
# after DeleteBridgePort("host0-3") in order A (main.go:441-460: three loose deletes)
gone = vf_name(0, 3)
c = [r for r in a if f"in_port={gone}," not in r and f"dl_dst={vf_macs[3]}," not in r]
fixture = dict(
    source="examples/*.yaml + marvell/ovs-dp/ovsdp.go format strings + marvell/main.go call order",
    network_functions=nfs, num_vfs=num_vfs, rpm_port=RPM, vf_macs=vf_macs, nf_macs=nf_macs, dp_interfaces=dp_ifs,
    formats=dict(hairpin=HAIRPIN, mac=MACRULE, wildcard=WILDCARD),
    rules_ports_then_nf=a, rules_nf_then_ports=b, rules_after_delete_host0_3=c, deleted_bridge_port="host0-3",
    ports=[RPM] + [vf_name(0, i) for i in range(num_vfs)] + dp_ifs,
)
assert len(a) == 3 * num_vfs + 2 == 26, len(a)
path = os.path.join(os.path.dirname(__file__), "replay_rules.json")
json.dump(fixture, open(path, "w"), indent=1)
print(f"wrote {path}: {len(a)} rules, NFs {nfs}")
