// mrvl_vsp.cpp — see mrvl_vsp.h.  Line references are to marvell/main.go of the reference.
#include "mrvl_vsp.h"

#include <errno.h>

#include <cstdio>

namespace sfc {

namespace {
// Everything one RPC programs becomes visible with one table commit (B200DP::BeginBatch / EndBatch).
struct RpcBatch {
    B200DP& dp;
    explicit RpcBatch(B200DP& d) : dp(d) { dp.BeginBatch(); }
    // rc = what the RPC body returned (its error text is already in place); a failing commit only shows
    // when the body itself succeeded
    int end(int rc, bool* commit_failed) {
        const int rc2 = dp.EndBatch();
        done = true;
        *commit_failed = !rc && rc2;
        return rc ? rc : rc2;
    }
    ~RpcBatch() {
        if (!done) dp.EndBatch();
    }
    bool done = false;
};
}  // namespace

int MrvlVsp::fail(int rc) {
    if (rc) err_ = dp_.LastError();
    return rc;
}

// doInit (:241-293): on the DPU side create the bridge; the NF interfaces exist by then
int MrvlVsp::Init(bool dpuMode) {
    is_dpu_mode_ = dpuMode;
    if (!dpuMode) return 0;  // host side: nothing to program
    return fail(dp_.InitDataPlane(kBridgeName));
}

// getVFDetails (:306-343): "host<pf>-<vf>" -> VF name; VF id 0 is the PF, hence the +1 (:321)
int MrvlVsp::getVFDetails(const std::string& name, std::string* vfName) {
    int pf = 0, vf = 0;
    char tail;
    if (sscanf(name.c_str(), "host%d-%d%c", &pf, &vf, &tail) != 2) {
        err_ = "no VFId Match Found";  // the reference's error text (:311)
        return -EINVAL;
    }
    char buf[32];
    snprintf(buf, sizeof(buf), "vf%d-%d", pf, vf + 1);
    *vfName = buf;
    return 0;
}

// CreateBridgePort (:347-421)
int MrvlVsp::CreateBridgePort(const std::string& name, const uint8_t mac[6], std::string* out_name) {
    RpcBatch batch(dp_);
    bool cf;
    const int rc = batch.end(createBridgePort(name, mac, out_name), &cf);
    return cf ? fail(rc) : rc;
}
int MrvlVsp::createBridgePort(const std::string& name, const uint8_t mac[6], std::string* out_name) {
    std::string vfName;
    int rc = getVFDetails(name, &vfName);
    if (rc) return rc;
    if ((rc = dp_.AddPortToDataPlane(kBridgeName, vfName, "", false))) return fail(rc);
    const std::string macs = mac_to_string(mac);
    network_store_[kNfName].vfPort.push_back(VfInfo{vfName, macs});  // (:361-381) created empty if absent
    if (is_nf_) {  // (:383-403) three rules per VF once an NF exists
        const NfPortMap& nf = network_store_[kNfName];
        if ((rc = dp_.AddFlowRuleToDataPlane(kBridgeName, vfName, nf.inpPort, ""))) return fail(rc);
        if ((rc = dp_.AddFlowRuleToDataPlane(kBridgeName, nf.inpPort, vfName, macs))) return fail(rc);
        if ((rc = dp_.AddFlowRuleToDataPlane(kBridgeName, nf.outPort, nf.outPort, macs))) return fail(rc);
    }
    if (out_name) *out_name = "bridge_port/" + name;
    return 0;
}

// DeleteBridgePort (:425-486)
int MrvlVsp::DeleteBridgePort(const std::string& name) {
    RpcBatch batch(dp_);
    bool cf;
    const int rc = batch.end(deleteBridgePort(name), &cf);
    return cf ? fail(rc) : rc;
}
int MrvlVsp::deleteBridgePort(const std::string& name) {
    std::string vfName;
    int rc = getVFDetails(name, &vfName);
    if (rc) return rc;
    NfPortMap& nf = network_store_[kNfName];
    if (is_nf_) {
        std::string vf_mac;
        for (auto& vf : nf.vfPort)
            if (vf.vfName == vfName) {
                vf_mac = vf.mac;
                break;
            }
        if ((rc = dp_.DeleteFlowRuleFromDataPlane(kBridgeName, vfName, "", ""))) return fail(rc);
        // the reference swallows errors of the next two deletes (`return nil, nil`, :452,:458)
        if (dp_.DeleteFlowRuleFromDataPlane(kBridgeName, nf.outPort, nf.outPort, vf_mac)) return 0;
        if (dp_.DeleteFlowRuleFromDataPlane(kBridgeName, nf.inpPort, vfName, vf_mac)) return 0;
    }
    if ((rc = dp_.DeletePortFromDataPlane(kBridgeName, vfName))) return fail(rc);
    for (size_t i = 0; i < nf.vfPort.size(); ++i)
        if (nf.vfPort[i].vfName == vfName) {
            nf.vfPort.erase(nf.vfPort.begin() + i);
            break;
        }
    return 0;
}

// CreateNetworkFunction (:490-497): MACs of the NF-side interfaces -> dataplane-side names
int MrvlVsp::CreateNetworkFunction(const std::string& input, const std::string& output) {
    is_nf_ = true;
    RpcBatch batch(dp_);
    // an unknown MAC yields the zero value "" in Go (:493-494) and the add-port then fails
    bool cf;
    const int rc = batch.end(AddNetworkFunction(device_store_[input], device_store_[output], kNfName), &cf);
    return cf ? fail(rc) : rc;
}

// AddNetworkFunction (:501-563)
int MrvlVsp::AddNetworkFunction(const std::string& inp, const std::string& out, const std::string& nfName) {
    int rc;
    if ((rc = dp_.AddPortToDataPlane(kBridgeName, inp, "", false))) return fail(rc);
    if ((rc = dp_.AddPortToDataPlane(kBridgeName, out, "", false))) return fail(rc);
    auto it = network_store_.find(nfName);
    if (it != network_store_.end()) {
        it->second.inpPort = inp;
        it->second.outPort = out;
        for (auto& vf : it->second.vfPort) {  // (:519-536)
            if ((rc = dp_.AddFlowRuleToDataPlane(kBridgeName, vf.vfName, inp, ""))) return fail(rc);
            if ((rc = dp_.AddFlowRuleToDataPlane(kBridgeName, inp, vf.vfName, vf.mac))) return fail(rc);
            if ((rc = dp_.AddFlowRuleToDataPlane(kBridgeName, out, out, vf.mac))) return fail(rc);
        }
    } else {
        NfPortMap m;
        m.inpPort = inp;
        m.outPort = out;
        network_store_[nfName] = m;
    }
    // out port <-> RPM (wire) interface (:540-560)
    const std::string& rpm = dp_.rpm_port();
    if ((rc = dp_.AddFlowRuleToDataPlane(kBridgeName, out, rpm, ""))) return fail(rc);
    if ((rc = dp_.AddFlowRuleToDataPlane(kBridgeName, rpm, out, ""))) return fail(rc);
    return 0;
}

// DeleteNetworkFunction (:565-572)
int MrvlVsp::DeleteNetworkFunction(const std::string& input, const std::string& output) {
    is_nf_ = false;
    RpcBatch batch(dp_);
    bool cf;
    const int rc = batch.end(DeleteNetworkFunctionPort(device_store_[input], device_store_[output], kNfName), &cf);
    return cf ? fail(rc) : rc;
}

// DeleteNetworkFunctionPort (:575-618)
int MrvlVsp::DeleteNetworkFunctionPort(const std::string& inp, const std::string& out, const std::string& nfName) {
    int rc;
    for (auto& vf : network_store_[nfName].vfPort)
        if ((rc = dp_.DeleteFlowRuleFromDataPlane(kBridgeName, vf.vfName, inp, ""))) return fail(rc);
    if ((rc = dp_.DeleteFlowRuleFromDataPlane(kBridgeName, inp, "", ""))) return fail(rc);
    if ((rc = dp_.DeleteFlowRuleFromDataPlane(kBridgeName, out, "", ""))) return fail(rc);
    if ((rc = dp_.DeleteFlowRuleFromDataPlane(kBridgeName, dp_.rpm_port(), "", ""))) return fail(rc);
    if ((rc = dp_.DeletePortFromDataPlane(kBridgeName, inp))) return fail(rc);
    if ((rc = dp_.DeletePortFromDataPlane(kBridgeName, out))) return fail(rc);
    return 0;
}

}  // namespace sfc
