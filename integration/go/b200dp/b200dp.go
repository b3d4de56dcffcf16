// Package b200dp is the third implementation of the Marvell VSP's mrvldp interface
// (internal/daemon/vendor-specific-plugins/marvell/main.go:54-62 of dpu-operator), next to ovs-dp and
// debug-dp: it forwards each call to libb200sfc.so through cgo.
//
// NOT COMPILED in this repository: the build image has no Go toolchain.  The same logic, method for
// method, is exercised in C++ (dpu-operator_b200/csrc/host/b200dp.cpp) by tests/test_replay.py and
// tests/test_vsp_grpc.py.  To use it: copy this directory to
// internal/daemon/vendor-specific-plugins/marvell/b200-dp/, put include/b200sfc.h on the cgo include
// path and libb200sfc.so on the linker path, build the VSP with CGO_ENABLED=1, and select it in
// NewMarvellVspServer (main.go:808-814) as INTEGRATION.md section 1 shows.
package b200dp

/*
#cgo LDFLAGS: -lb200sfc
#include <stdlib.h>
#include "b200sfc.h"
*/
import "C"
import ("errors"; "net"; "strings"; "unsafe")

// batch > 0: the seven methods stage only; EndBatch makes everything visible with one table commit
// (one commit per vendor-plugin RPC, as dpu-operator_b200/csrc/host/b200dp.cpp does).
type B200DP struct {
	h       *C.sfc_handle
	batch   int
	pending bool
}

func NewB200DP() *B200DP { return &B200DP{} }

func (d *B200DP) err(rc C.int) error {
	if rc == 0 { return nil }
	return errors.New(C.GoString(C.sfc_last_error(d.h)))
}
func (d *B200DP) commit() error {
	if d.batch > 0 { d.pending = true; return nil }
	return d.err(C.sfc_commit(d.h))
}
func (d *B200DP) BeginBatch() { d.batch++ }
func (d *B200DP) EndBatch() error {
	if d.batch > 0 { d.batch-- }
	if d.batch > 0 || !d.pending { return nil }
	d.pending = false
	return d.err(C.sfc_commit(d.h))
}
func (d *B200DP) port(name string) (C.uint16_t, error) {
	cs := C.CString(name); defer C.free(unsafe.Pointer(cs))
	var id C.uint16_t
	if rc := C.sfc_port_lookup(d.h, cs, &id); rc != 0 { return 0, errors.New("unknown port " + name) }
	return id, nil
}
func (d *B200DP) InitDataPlane(bridgeName string) error {
	if rc := C.sfc_create(0, &d.h); rc != 0 { return errors.New("sfc_create failed: no B200") }
	return d.err(C.sfc_commit(d.h))
}
func (d *B200DP) AddPortToDataPlane(bridgeName, portName, vfPCIAddres string, isDPDK bool) error {
	cs := C.CString(portName); defer C.free(unsafe.Pointer(cs))
	if rc := C.sfc_port_add(d.h, cs, nil); rc != 0 { return d.err(rc) }
	return d.commit()
}
func (d *B200DP) DeletePortFromDataPlane(bridgeName, portName string) error {
	cs := C.CString(portName); defer C.free(unsafe.Pointer(cs))
	if rc := C.sfc_port_del(d.h, cs); rc != 0 { return d.err(rc) }
	return d.commit()
}
func (d *B200DP) ReadAllPortFromDataPlane(bridgeName string) (string, error) {
	n := C.sfc_port_list(d.h, nil, 0)
	buf := (*C.char)(C.malloc(C.size_t(n))); defer C.free(unsafe.Pointer(buf))
	C.sfc_port_list(d.h, buf, C.size_t(n))
	return C.GoString(buf), nil
}
func (d *B200DP) DeleteDataplane(bridgeName string) error { return d.err(C.sfc_destroy(d.h)) }
func (d *B200DP) AddFlowRuleToDataPlane(bridgeName, inPort, outPort, dstMac string) error {
	in, err := d.port(inPort); if err != nil { return err }
	out, err := d.port(outPort); if err != nil { return err }
	var mac *C.uint8_t
	if dstMac != "" {
		hw, err := net.ParseMAC(dstMac); if err != nil { return err }
		mac = (*C.uint8_t)(C.CBytes(hw)); defer C.free(unsafe.Pointer(mac))
	}
	if rc := C.sfc_l2_rule_add(d.h, in, out, mac, nil); rc != 0 { return d.err(rc) }
	return d.commit()
}
func (d *B200DP) DeleteFlowRuleFromDataPlane(bridgeName, inPort, outPort, dstMac string) error {
	in, err := d.port(inPort); if err != nil { return err }
	var mac *C.uint8_t
	if dstMac != "" {
		hw, err := net.ParseMAC(dstMac); if err != nil { return err }
		mac = (*C.uint8_t)(C.CBytes(hw)); defer C.free(unsafe.Pointer(mac))
	}
	if rc := C.sfc_l2_rule_del(d.h, in, mac); rc != 0 { return d.err(rc) }
	return d.commit()
}
var _ = strings.TrimSpace
