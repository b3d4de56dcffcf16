// vsp_capi.cpp — extern "C" wrappers of include/b200vsp.h over MrvlVsp.
#include <errno.h>

#include <algorithm>
#include <cstring>

#include "../../../include/b200vsp.h"
#include "mrvl_vsp.h"

struct vsp_server {
    sfc::MrvlVsp vsp;
    std::string err;
    vsp_server(int device, const char* rpm) : vsp(device, rpm && *rpm ? rpm : "rpm0") {}
};

static int done(vsp_server* v, int rc) {
    if (rc) v->err = v->vsp.LastError();
    return rc;
}

extern "C" {
int vsp_create(int device, const char* rpm_port_name, vsp_server** out) {
    if (!out) return -EINVAL;
    *out = new vsp_server(device, rpm_port_name);
    return 0;
}
int vsp_destroy(vsp_server* v) {
    delete v;
    return 0;
}
const char* vsp_last_error(const vsp_server* v) { return v ? v->err.c_str() : "null server"; }
int vsp_add_nf_device(vsp_server* v, const char* nf_mac, const char* dp_interface) {
    if (!v || !nf_mac || !dp_interface) return -EINVAL;
    v->vsp.AddNfDevice(nf_mac, dp_interface);
    return 0;
}
int vsp_init(vsp_server* v, int dpu_mode) { return v ? done(v, v->vsp.Init(dpu_mode != 0)) : -EINVAL; }
int vsp_create_bridge_port(vsp_server* v, const char* name, const uint8_t mac[6], char* out_name, size_t out_len) {
    if (!v || !name || !mac) return -EINVAL;
    std::string out;
    int rc = done(v, v->vsp.CreateBridgePort(name, mac, &out));
    if (!rc && out_name && out_len) {
        strncpy(out_name, out.c_str(), out_len - 1);
        out_name[out_len - 1] = 0;
    }
    return rc;
}
int vsp_delete_bridge_port(vsp_server* v, const char* name) {
    return v && name ? done(v, v->vsp.DeleteBridgePort(name)) : -EINVAL;
}
int vsp_create_network_function(vsp_server* v, const char* input, const char* output) {
    return v && input && output ? done(v, v->vsp.CreateNetworkFunction(input, output)) : -EINVAL;
}
int vsp_delete_network_function(vsp_server* v, const char* input, const char* output) {
    return v && input && output ? done(v, v->vsp.DeleteNetworkFunction(input, output)) : -EINVAL;
}
long vsp_flow_dump(const vsp_server* v, char* buf, size_t len) {
    if (!v) return -EINVAL;
    std::string s;
    for (auto& line : const_cast<vsp_server*>(v)->vsp.dataplane().FlowDump()) s += line + "\n";
    if (buf && len) {
        size_t c = std::min(len - 1, s.size());
        memcpy(buf, s.data(), c);
        buf[c] = 0;
    }
    return (long)s.size() + 1;
}
sfc_handle* vsp_dataplane(vsp_server* v) { return v ? v->vsp.dataplane().handle() : nullptr; }
}
