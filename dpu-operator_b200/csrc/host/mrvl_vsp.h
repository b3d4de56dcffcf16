// mrvl_vsp.h — the Marvell vendor plugin's rule logic, restated over the B200 dataplane.
//
// Reference: mrvlVspServer in internal/daemon/vendor-specific-plugins/marvell/main.go —
// Init (:241-303), CreateBridgePort (:347-421), DeleteBridgePort (:425-486), CreateNetworkFunction /
// AddNetworkFunction (:490-563), DeleteNetworkFunction / DeleteNetworkFunctionPort (:565-618).
// Method names, argument meaning, the order in which ports and rules are programmed, and the
// error behaviour follow the reference.  What cannot exist here is replaced by a stated stand-in:
//   - the veth pairs created by ConfigureNetworkInterface (:183-215) and their kernel-assigned
//     MACs -> AddNfDevice(nfMac, dpInterfaceName);
//   - VF netdev names looked up in sysfs (getVFDetails :306-343) -> the reference's own DPDK-mode
//     naming "vf<pf>-<vf+1>" (:333);
//   - the RPM (wire) interface found by PCI device id a063 (:540) -> a configured name.
#pragma once
#include <map>
#include <string>
#include <vector>

#include "b200dp.h"

namespace sfc {

class MrvlVsp {
  public:
    static constexpr const char* kNfName = "mrvl-nf1";     // main.go:49
    static constexpr const char* kBridgeName = "br-mrv0";  // main.go:272
    static constexpr int kNoOfPortPairs = 2;               // main.go:40

    explicit MrvlVsp(int device, std::string rpm_port = "rpm0") : dp_(device, std::move(rpm_port)) {}

    // stand-in for createVethPair (:105-152): deviceStore[nfMac] = {dpInterfaceName}
    void AddNfDevice(const std::string& nfMac, const std::string& dpInterfaceName) { device_store_[nfMac] = dpInterfaceName; }

    int Init(bool dpuMode);
    // out_name receives "bridge_port/<name>" (main.go:415)
    int CreateBridgePort(const std::string& name, const uint8_t mac[6], std::string* out_name);
    int DeleteBridgePort(const std::string& name);
    int CreateNetworkFunction(const std::string& input, const std::string& output);
    int DeleteNetworkFunction(const std::string& input, const std::string& output);

    const std::string& LastError() const { return err_; }
    B200DP& dataplane() { return dp_; }

  private:
    struct VfInfo {
        std::string vfName, mac;
    };
    struct NfPortMap {  // mrvlNfPortMap
        std::vector<VfInfo> vfPort;
        std::string inpPort, outPort;
    };
    int fail(int rc);
    int createBridgePort(const std::string& name, const uint8_t mac[6], std::string* out_name);  // RPC bodies: one commit per RPC around them
    int deleteBridgePort(const std::string& name);
    int getVFDetails(const std::string& bridgePortName, std::string* vfName);
    int AddNetworkFunction(const std::string& inp, const std::string& out, const std::string& nfName);
    int DeleteNetworkFunctionPort(const std::string& inp, const std::string& out, const std::string& nfName);

    B200DP dp_;
    std::string err_;
    std::map<std::string, std::string> device_store_;  // NF-side MAC -> dataplane-side interface
    std::map<std::string, NfPortMap> network_store_;
    bool is_nf_ = false;
    bool is_dpu_mode_ = false;
};

}  // namespace sfc
