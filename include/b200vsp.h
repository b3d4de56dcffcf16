/* b200vsp.h — C view of the host-side mirror of the reference's Marvell vendor plugin
 * (dpu-operator_b200/csrc/host/), so that the same logic a Go VSP would run can be driven from C,
 * ctypes or a gRPC front door.  Exported by libb200sfc.so.
 *
 * Each call is the RPC handler of the same name in
 * internal/daemon/vendor-specific-plugins/marvell/main.go (cited per function); the VSP programs
 * its rules through the `mrvldp` mirror (B200DP), i.e. through include/b200sfc.h.
 * Return: 0 or a negative errno; vsp_last_error() gives the text of the Go `error`. */
#ifndef B200VSP_H
#define B200VSP_H
#include <stddef.h>
#include <stdint.h>

#include "b200sfc.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vsp_server vsp_server;

/* NewMarvellVspServer (main.go:795-822).  device -1 = offline (rule staging only). */
int vsp_create(int device, const char* rpm_port_name, vsp_server** out);
int vsp_destroy(vsp_server* v);
const char* vsp_last_error(const vsp_server* v);

/* Stand-in for the veth pairs of ConfigureNetworkInterface (main.go:105-152,183-215): declares
 * that the NF-side interface with MAC `nf_mac` ("aa:bb:..") has dataplane-side peer `dp_interface`. */
int vsp_add_nf_device(vsp_server* v, const char* nf_mac, const char* dp_interface);

/* LifeCycleService.Init (main.go:297-303; dpu-api/api.proto:7-9) */
int vsp_init(vsp_server* v, int dpu_mode);
/* BridgePortService.CreateBridgePort / DeleteBridgePort (main.go:347-421, 425-486).
 * name is "host<pf>-<vf>" (internal/daemon/hostsidemanager.go:59-71); out_name receives
 * "bridge_port/<name>". */
int vsp_create_bridge_port(vsp_server* v, const char* name, const uint8_t mac[6], char* out_name, size_t out_len);
int vsp_delete_bridge_port(vsp_server* v, const char* name);
/* NetworkFunctionService.Create / DeleteNetworkFunction (main.go:490-497, 565-572;
 * dpu-api/api.proto:11-14): input/output are the MACs the daemon paired
 * (internal/daemon/dpusidemanager.go:104-119). */
int vsp_create_network_function(vsp_server* v, const char* input, const char* output);
int vsp_delete_network_function(vsp_server* v, const char* input, const char* output);

/* The live steering rules, one per line, spelled as ovs-dp/ovsdp.go:120,127,135 would have passed
 * them to `ovs-ofctl add-flow`.  Returns bytes needed including NUL. */
long vsp_flow_dump(const vsp_server* v, char* buf, size_t len);
/* The dataplane handle the rules were programmed into (for sfc_classify etc.); NULL before Init. */
sfc_handle* vsp_dataplane(vsp_server* v);

#ifdef __cplusplus
}
#endif
#endif
