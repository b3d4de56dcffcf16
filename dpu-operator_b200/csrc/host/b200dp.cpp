// b200dp.cpp — see b200dp.h.  Each method says which OVS command of the reference backend it stands for.
#include "b200dp.h"

#include <errno.h>

#include <algorithm>
#include <cstdio>

namespace sfc {

bool parse_mac(const std::string& s, uint8_t out[6]) {
    unsigned v[6];
    char tail;
    if (sscanf(s.c_str(), "%2x:%2x:%2x:%2x:%2x:%2x%c", &v[0], &v[1], &v[2], &v[3], &v[4], &v[5], &tail) != 6) return false;
    for (int i = 0; i < 6; ++i) out[i] = (uint8_t)v[i];
    return true;
}
std::string mac_to_string(const uint8_t mac[6]) {  // net.HardwareAddr.String(): lower-case, colon separated
    char buf[18];
    snprintf(buf, sizeof(buf), "%02x:%02x:%02x:%02x:%02x:%02x", mac[0], mac[1], mac[2], mac[3], mac[4], mac[5]);
    return buf;
}

B200DP::~B200DP() {
    if (h_) sfc_destroy(h_);
}

int B200DP::fail(int rc) {
    if (rc) err_ = h_ ? sfc_last_error(h_) : "dataplane not initialised";
    return rc;
}

int B200DP::commit() {
    if (batch_depth_ > 0) {
        commit_pending_ = true;
        return 0;
    }
    return sfc_commit(h_);
}

int B200DP::EndBatch() {
    if (batch_depth_ > 0 && --batch_depth_ > 0) return 0;
    if (!commit_pending_ || !h_) return 0;
    commit_pending_ = false;
    return fail(sfc_commit(h_));
}

// ovs-vsctl --may-exist add-br <bridge> + add the RPM (wire) port — ovs-dp/ovsdp.go:71-102
int B200DP::InitDataPlane(const std::string& bridgeName) {
    if (h_) return 0;  // --may-exist
    int rc = sfc_create(device_, &h_);
    if (rc) {
        err_ = "sfc_create failed: no B200 on this host (there is no CPU fallback)";
        return rc;
    }
    bridge_ = bridgeName;
    // the reference logs and carries on if the RPM port cannot be added (ovsdp.go:91-99); so do we
    AddPortToDataPlane(bridgeName, rpm_port_, "", false);
    return fail(sfc_commit(h_));
}

// ovs-vsctl --may-exist add-port <bridge> <port> — ovs-dp/ovsdp.go:40-55
int B200DP::AddPortToDataPlane(const std::string&, const std::string& portName, const std::string&, bool) {
    if (!h_) return fail(-ENODEV);
    int rc = sfc_port_add(h_, portName.c_str(), nullptr);
    return rc ? fail(rc) : fail(commit());
}

// ovs-vsctl del-port <bridge> <port> — ovs-dp/ovsdp.go:58-62
int B200DP::DeletePortFromDataPlane(const std::string&, const std::string& portName) {
    if (!h_) return fail(-ENODEV);
    int rc = sfc_port_del(h_, portName.c_str());  // also drops the steering rules that name the port
    if (rc) return fail(rc);
    flows_.erase(std::remove_if(flows_.begin(), flows_.end(), [&](const Flow& f) { return f.in == portName || f.out == portName; }),
                 flows_.end());
    return fail(commit());
}

// ovs-vsctl list-ports <bridge> — ovs-dp/ovsdp.go:104-112
int B200DP::ReadAllPortFromDataPlane(const std::string&, std::string* out) {
    if (!h_) return fail(-ENODEV);
    long n = sfc_port_list(h_, nullptr, 0);
    std::string buf((size_t)n, '\0');
    sfc_port_list(h_, &buf[0], (size_t)n);
    buf.resize(n ? n - 1 : 0);
    if (out) *out = buf;
    return 0;
}

// ovs-vsctl del-br <bridge> — ovs-dp/ovsdp.go:66-69
int B200DP::DeleteDataplane(const std::string&) {
    if (!h_) return 0;
    int rc = sfc_destroy(h_);
    h_ = nullptr;
    flows_.clear();
    return rc;
}

// ovs-ofctl add-flow — ovs-dp/ovsdp.go:113-142.  The rule shape (and its OpenFlow priority) follows
// from (dstMac empty?, inPort == outPort?) exactly as in the reference; the library applies it.
int B200DP::AddFlowRuleToDataPlane(const std::string&, const std::string& inPort, const std::string& outPort,
                                   const std::string& dstMac) {
    if (!h_) return fail(-ENODEV);
    uint16_t in = 0, out = 0;
    if (sfc_port_lookup(h_, inPort.c_str(), &in) || sfc_port_lookup(h_, outPort.c_str(), &out)) {
        // ovs-ofctl fails on an unknown port name; so does this backend
        err_ = "add-flow: unknown port '" + (sfc_port_lookup(h_, inPort.c_str(), &in) ? inPort : outPort) + "'";
        return -ENOENT;
    }
    uint8_t mac[6];
    if (!dstMac.empty() && !parse_mac(dstMac, mac)) {
        err_ = "add-flow: bad MAC '" + dstMac + "'";
        return -EINVAL;
    }
    int rc = sfc_l2_rule_add(h_, in, out, dstMac.empty() ? nullptr : mac, nullptr);
    if (rc) return fail(rc);
    // an identical match replaces the actions (ovs-ofctl add-flow semantics)
    auto shape = [](const std::string& in, const std::string& out, const std::string& mac) {
        return mac.empty() ? 0 : (in == out ? 1 : 2);  // wildcard / hairpin / MAC-qualified: distinct priorities
    };
    auto same = [&](const Flow& f) {
        return f.in == inPort && f.mac == dstMac && shape(f.in, f.out, f.mac) == shape(inPort, outPort, dstMac);
    };
    auto it = std::find_if(flows_.begin(), flows_.end(), same);
    if (it != flows_.end()) it->out = outPort;
    else flows_.push_back(Flow{inPort, outPort, dstMac});
    return fail(commit());
}

// ovs-ofctl del-flows <bridge> in_port=S[,dl_dst=M] — ovs-dp/ovsdp.go:143-162 (a loose match)
int B200DP::DeleteFlowRuleFromDataPlane(const std::string&, const std::string& inPort, const std::string&,
                                        const std::string& dstMac) {
    if (!h_) return fail(-ENODEV);
    uint16_t in = 0;
    if (sfc_port_lookup(h_, inPort.c_str(), &in)) {
        err_ = "del-flows: unknown port '" + inPort + "'";
        return -ENOENT;
    }
    uint8_t mac[6];
    if (!dstMac.empty() && !parse_mac(dstMac, mac)) {
        err_ = "del-flows: bad MAC '" + dstMac + "'";
        return -EINVAL;
    }
    int rc = sfc_l2_rule_del(h_, in, dstMac.empty() ? nullptr : mac);
    if (rc) return fail(rc);
    flows_.erase(std::remove_if(flows_.begin(), flows_.end(),
                                [&](const Flow& f) { return f.in == inPort && (dstMac.empty() || f.mac == dstMac); }),
                 flows_.end());
    return fail(commit());
}

std::vector<std::string> B200DP::FlowDump() const {
    std::vector<std::string> out;
    for (auto& f : flows_) {
        if (f.mac.empty())
            out.push_back("priority=10,in_port=" + f.in + ",actions=output:" + f.out);
        else if (f.in == f.out)
            out.push_back("priority=100,in_port=" + f.in + ",dl_dst=" + f.mac + ",actions=in_port");
        else
            out.push_back("in_port=" + f.in + ",dl_dst=" + f.mac + ",actions=output:" + f.out);
    }
    return out;
}

}  // namespace sfc
