// b200dp.h — the B200 implementation of the reference's in-process dataplane interface.
//
// Reference: `type mrvldp interface` — internal/daemon/vendor-specific-plugins/marvell/main.go:54-62,
// implemented there by ovsdp.OvsDP (marvell/ovs-dp/ovsdp.go) and debugdp.DebugDP
// (marvell/debug-dp/debugdp.go).  The seven methods keep their reference names and argument order;
// Go's `error` becomes an int (0 / negative errno) plus LastError().  This is C++ only because no
// Go toolchain exists in the build image; INTEGRATION.md holds the cgo version.
#pragma once
#include <string>
#include <vector>

#include "../../../include/b200sfc.h"

namespace sfc {

class B200DP {
  public:
    // device -1 = offline handle (rule staging and image building only; CPU tests)
    explicit B200DP(int device, std::string rpm_port = "rpm0") : device_(device), rpm_port_(std::move(rpm_port)) {}
    ~B200DP();
    B200DP(const B200DP&) = delete;
    B200DP& operator=(const B200DP&) = delete;

    int InitDataPlane(const std::string& bridgeName);
    int AddPortToDataPlane(const std::string& bridgeName, const std::string& portName, const std::string& vfPCIAddres,
                           bool isDPDK);
    int DeletePortFromDataPlane(const std::string& bridgeName, const std::string& portName);
    int ReadAllPortFromDataPlane(const std::string& bridgeName, std::string* out);
    int DeleteDataplane(const std::string& bridgeName);
    int AddFlowRuleToDataPlane(const std::string& bridgeName, const std::string& inPort, const std::string& outPort,
                               const std::string& dstMac);
    int DeleteFlowRuleFromDataPlane(const std::string& bridgeName, const std::string& inPort, const std::string& outPort,
                                    const std::string& dstMac);

    // One RPC of the vendor plugin programs several ports and rules (CreateBridgePort: one port + three rules,
    // marvell/main.go:347-421).  Between BeginBatch and EndBatch the methods above only stage; EndBatch makes
    // everything staged visible with ONE table commit.  Outside a batch every method commits by itself, like
    // one ovs-vsctl / ovs-ofctl invocation of the reference backend.
    void BeginBatch() { ++batch_depth_; }
    int EndBatch();

    const std::string& LastError() const { return err_; }
    sfc_handle* handle() const { return h_; }
    const std::string& rpm_port() const { return rpm_port_; }
    // The live rule set, one line per rule, spelled as the OVS backend would have passed it to
    // `ovs-ofctl add-flow` (ovs-dp/ovsdp.go:120,127,135) — what the replay test compares.
    std::vector<std::string> FlowDump() const;

  private:
    struct Flow {
        std::string in, out, mac;
    };
    int fail(int rc);
    int commit();
    int batch_depth_ = 0;
    bool commit_pending_ = false;
    int device_;
    std::string rpm_port_, bridge_, err_;
    sfc_handle* h_ = nullptr;
    std::vector<Flow> flows_;
};

bool parse_mac(const std::string& s, uint8_t out[6]);
std::string mac_to_string(const uint8_t mac[6]);

}  // namespace sfc
