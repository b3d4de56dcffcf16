"""Regenerate tests/golden/intel_rules.json — the rule stream of the Intel IPU vendor plugin for "8 host VFs + 1 NF",
computed FROM THE GO SOURCE of the vendored plugin (run in the build container, where /root/reference is mounted).

Nothing is retyped: for each rule builder of vendor/github.com/intel/ipu-opi-plugins/ipu-plugin/pkg/p4rtclient/
p4rtclient.go this script cuts the function body out of the file, reads every `Control:` / `Metadata: fmt.Sprintf(
"<format>", <args>)` pair of its add-entry rule set with a regex, reads which Go variable is the VSI / vport of
which parameter from the `x, y, err := utils.ExtractVsiVportInfo(<param>)` lines, and evaluates the Sprintf calls
with `VSI = second MAC octet`, `vport = VSI + vsiToVportOffset` (constant read from pkg/utils/utils.go).  Only the
call order of AddNFP4Rules (:819-859), AddPeerToPeerP4Rules (:903-919) and the two RPCs is transliterated below.

tests/test_intel_vsp.py holds dpu-operator_b200/intel_vsp.py to this fixture string for string.
    python tests/golden/make_intel_rules.py
"""
import json
import os
import re

PKG = "/root/reference/vendor/github.com/intel/ipu-opi-plugins/ipu-plugin/pkg"
SRC = open(os.path.join(PKG, "p4rtclient/p4rtclient.go")).read()
OFFSET = int(re.search(r"vsiToVportOffset\s*=\s*(\d+)", open(os.path.join(PKG, "utils/utils.go")).read()).group(1))
assert re.search(r"vfVsi := int\(macAddrByte\[1\]\)", open(os.path.join(PKG, "utils/utils.go")).read())  # VSI = mac[1]


def body(func):
    m = re.search(r"\nfunc " + func + r"\((.*?)\) error \{\n(.*?)\n\}\n", SRC, re.S)
    params = [p.strip().split()[0] for p in m.group(1).split(",")][1:]  # drop p4rtClient
    return params, m.group(2)


def builder(func):
    """-> (params, [(table, format, [arg variable names])], {variable: (param, 'vsi'|'vport')})"""
    params, text = body(func)
    binds = {}
    for vsi_var, vport_var, arg in re.findall(r"(\w+), (\w+), err :?= utils\.ExtractVsiVportInfo\(([\w.()]+)\)", text):
        arg = {"hostMacAddr.String()": "hostVfMac"}.get(arg, arg)  # AddHostVfP4Rules wraps its []byte parameter (:648)
        assert arg in params, (func, arg, params)
        binds[vsi_var], binds[vport_var] = (arg, "vsi"), (arg, "vport")
    rules = []
    for action, control, fmt, args in re.findall(
            r'Action:\s+"([\w-]+)",\s+P4br:\s+"br0",\s+Control:\s+"([\w.]+)",\s+Metadata: fmt\.Sprintf\(\s*"([^"]+)",\s*([^)]*?),?\s*\),', text):
        assert action == "add-entry", (func, action)
        rules.append((control, fmt, [a.strip() for a in args.split(",") if a.strip()]))
    return params, rules, binds


def run(func, **macs):
    params, rules, binds = builder(func)
    out = []
    for control, fmt, args in rules:
        vals = []
        for a in args:
            param, what = binds[a]
            vsi = bytes(int(x, 16) for x in macs[param].split(":"))[1]
            vals.append(vsi if what == "vsi" else vsi + OFFSET)
        out.append([control, fmt % tuple(vals)])
    return out


host_vf = lambda vf, acc: run("AddHostVfP4Rules", hostVfMac=vf, accMac=acc)  # noqa: E731
nf_pr = lambda a, b: run("programNfPrVportP4Rules", ingressMac=a, egressMac=b)  # noqa: E731
vsi_vsi = lambda a, b: run("programVsiToVsiP4Rules", mac1=a, mac2=b)  # noqa: E731
assert len(host_vf("00:01:00:00:03:14", "00:09:00:00:03:14")) == 5 and len(nf_pr("00:01:00:00:03:14", "00:09:00:00:03:14")) == 6


def add_nf(vf_macs, ingress, egress, ingress_pr, egress_pr):  # AddNFP4Rules :839-858
    out = nf_pr(ingress, ingress_pr) + nf_pr(egress, egress_pr)
    for vf in vf_macs:
        for nf in (ingress, egress):
            out += vsi_vsi(vf, nf)
    return out


def peer_to_peer(vf_macs):  # AddPeerToPeerP4Rules :912-916
    out = []
    for i in range(len(vf_macs)):
        for j in range(i + 1, len(vf_macs)):
            out += vsi_vsi(vf_macs[i], vf_macs[j])
    return out


# the scenario: 8 host VFs (VF MAC pattern 00:<vsi>:00:00:03:14, ipuplugin/bridgeport.go:99-100), 8 ACC port representors,
# one NF whose two ports have a port representor each
vf_macs = [f"00:{0x21 + i:02x}:00:00:03:14" for i in range(8)]
acc_macs = [f"00:{0x09 + i:02x}:00:00:03:14" for i in range(8)]
nf_macs = dict(ingress="00:31:00:00:03:14", egress="00:32:00:00:03:14", ingress_pr="00:11:00:00:03:14", egress_pr="00:12:00:00:03:14")
stream = []
for vf, acc in zip(vf_macs, acc_macs):  # CreateBridgePort x 8 (bridgeport.go:151)
    stream += host_vf(vf, acc)
stream += add_nf(vf_macs, nf_macs["ingress"], nf_macs["egress"], nf_macs["ingress_pr"], nf_macs["egress_pr"])  # CreateNetworkFunction
p2p = peer_to_peer(vf_macs)
# ---- the three builders that take their VSIs from the IMC: evaluated with the VSIs given, constants read from the file ----
CONSTS = {k: int(v) for k, v in re.findall(r"^\t(mirror_profile_id|bridgeId|phyPort)\s*=\s*(\d+)", SRC, re.M)}
assert set(CONSTS) == {"mirror_profile_id", "bridgeId", "phyPort"}


def run_imc(func, env):
    """Evaluate the add-entry Sprintf calls of `func` with the Go expressions in `env` (name -> value); literals pass through."""
    _, text = body(func)
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # rule sets kept "for reference only" inside comments
    out = []
    for control, meta in re.findall(r'Action:\s+"add-entry",\s+P4br:\s+"br0",\s+Control:\s+"([\w.]+)",\s+Metadata:\s+(.*?),\n\t+\}', text, re.S):
        m = re.match(r'fmt\.Sprintf\(\s*"([^"]+)",\s*(.*?),?\s*\)$', meta.strip(), re.S)
        if not m:
            out.append([control, meta.strip().strip('"')])
            continue
        fmt, args = m.group(1), [a.strip() for a in m.group(2).split(",") if a.strip()]
        vals = [eval(a, {}, dict(CONSTS, **env)) for a in args]  # the arguments are names or `name+16`
        out.append([control, fmt % tuple(vals)])
    return out


phy = run_imc("programPhyVportP4Rules", dict(phyPort=0, prVsi=0x0a, prVport=0x0a + OFFSET))
# AddLAGP4Rules: one literal entry, then `for idx := 0; idx < 8; idx++` around one Sprintf — the loop is unrolled here
assert re.search(r"for idx := 0; idx < 8; idx\+\+", body("AddLAGP4Rules")[1])
lag = []
for idx in range(8):
    both = run_imc("AddLAGP4Rules", dict(idx=idx))
    assert len(both) == 2
    lag = [both[0]] + lag[1:] + [both[1]]
rh = run_imc("AddRHPrimaryNetworkVportP4Rules", dict(d4Vsi=0x0c, d1Vsi=0x0d, d1MacAddr="000d00000314"))
assert len(phy) == 3 and len(lag) == 9 and len(rh) == 5, (len(phy), len(lag), len(rh))

fixture = dict(
    source="vendor/github.com/intel/ipu-opi-plugins/ipu-plugin/pkg/p4rtclient/p4rtclient.go: fmt.Sprintf calls of AddHostVfP4Rules, "
           "programNfPrVportP4Rules, programVsiToVsiP4Rules evaluated; call order of AddNFP4Rules / AddPeerToPeerP4Rules",
    vsi_to_vport_offset=OFFSET, vf_macs=vf_macs, acc_macs=acc_macs, nf_macs=nf_macs,
    formats={f: [[c, fmt, args] for c, fmt, args in builder(f)[1]] for f in ("AddHostVfP4Rules", "programNfPrVportP4Rules", "programVsiToVsiP4Rules")},
    rules_8_host_vfs_then_1_nf=stream, rules_peer_to_peer=p2p,
    consts=CONSTS, rules_phy_port_0_pr_vsi_10=phy, rules_lag=lag, rules_rh_primary_d4_vsi_12_d1_vsi_13=rh, rh_d1_mac="00:0d:00:00:03:14",
)
assert len(stream) == 8 * 5 + 2 * 6 + 8 * 2 * 2 == 84 and len(p2p) == 8 * 7
path = os.path.join(os.path.dirname(__file__), "intel_rules.json")
json.dump(fixture, open(path, "w"), indent=1)
print(f"wrote {path}: {len(stream)} + {len(p2p)} rules; vsiToVportOffset {OFFSET}")
