"""B200 vendor-specific plugin (VSP): the reference's own process-level plugin boundary.

The dpu-operator daemon talks to a VSP over gRPC on a unix socket
(``/var/run/dpu-daemon/vendor-plugin/vendor-plugin.sock``, internal/utils/path_manager.go:58-60).
A VSP registers four services on one server (marvell/main.go:738-742):

  /Vendor.LifeCycleService/Init
  /Vendor.NetworkFunctionService/{Create,Delete}NetworkFunction
  /Vendor.DeviceService/{GetDevices,SetNumVfs}                    (dpu-api/api.proto:7-36)
  /opi_api.network.evpn_gw.v1alpha1.BridgePortService/{Create,Delete}BridgePort
                                                  (vendored l2_xpu_infra_mgr_grpc.pb.go:284-288)

This module serves exactly those methods and translates them into the C++ mirror of the Marvell
VSP logic (``csrc/host/``, reached through ``include/b200vsp.h``), which programs the GPU
dataplane through the C ABI.  The daemon (``internal/daemon/plugin/vendorplugin.go:183-265``) needs no
change to use it: same socket, same methods, same messages.

No protoc is needed: the message types are declared below with ``descriptor_pb2`` — a restatement
of dpu-api/api.proto and of the OPI BridgePort messages with the field numbers of the vendored
``l2_xpu_infra_mgr.pb.go`` (:777-781, :844-849, :912-914, :975-977, :1200-1203).
"""
from __future__ import annotations

import threading
from concurrent import futures

import grpc
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

from .capi import SfcError, VspServer

VENDOR_PLUGIN_SOCKET = "/var/run/dpu-daemon/vendor-plugin/vendor-plugin.sock"  # path_manager.go:58-60
IPV6_ADDR_DPU, IPV6_ADDR_HOST, DEFAULT_PORT = "fe80::1", "fe80::2", 8085        # marvell/main.go:41-42,36
OPI = "opi_api.network.evpn_gw.v1alpha1"

_T = descriptor_pb2.FieldDescriptorProto


def _field(msg, name, number, ftype, label=_T.LABEL_OPTIONAL, type_name=None):
    f = msg.field.add()
    f.name, f.number, f.type, f.label = name, number, ftype, label
    if type_name:
        f.type_name = type_name
    return f


def _build_pool():
    pool = descriptor_pool.DescriptorPool()
    # ---- dpu-api/api.proto ----
    fd = descriptor_pb2.FileDescriptorProto(name="b200/dpu_api.proto", package="Vendor", syntax="proto3")
    m = fd.message_type.add(name="InitRequest")
    _field(m, "dpu_mode", 1, _T.TYPE_BOOL)
    _field(m, "dpu_identifier", 2, _T.TYPE_STRING)
    m = fd.message_type.add(name="IpPort")
    _field(m, "ip", 1, _T.TYPE_STRING)
    _field(m, "port", 2, _T.TYPE_INT32)
    m = fd.message_type.add(name="NFRequest")
    _field(m, "input", 1, _T.TYPE_STRING)
    _field(m, "output", 2, _T.TYPE_STRING)
    fd.message_type.add(name="Empty")
    m = fd.message_type.add(name="VfCount")
    _field(m, "vf_cnt", 1, _T.TYPE_INT32)
    m = fd.message_type.add(name="TopologyInfo")
    _field(m, "node", 1, _T.TYPE_STRING)
    m = fd.message_type.add(name="Device")
    _field(m, "ID", 1, _T.TYPE_STRING)
    _field(m, "health", 2, _T.TYPE_STRING)
    _field(m, "topology", 3, _T.TYPE_MESSAGE, type_name=".Vendor.TopologyInfo")
    m = fd.message_type.add(name="DeviceListResponse")
    e = m.nested_type.add(name="DevicesEntry")  # map<string, Device> devices = 1
    e.options.map_entry = True
    _field(e, "key", 1, _T.TYPE_STRING)
    _field(e, "value", 2, _T.TYPE_MESSAGE, type_name=".Vendor.Device")
    _field(m, "devices", 1, _T.TYPE_MESSAGE, _T.LABEL_REPEATED, ".Vendor.DeviceListResponse.DevicesEntry")
    pool.Add(fd)
    # ---- google.protobuf.Empty ----
    fe = descriptor_pb2.FileDescriptorProto(name="b200/empty.proto", package="google.protobuf", syntax="proto3")
    fe.message_type.add(name="Empty")
    try:
        pool.Add(fe)
    except TypeError:
        pass
    # ---- OPI evpn-gw BridgePort (subset the daemon uses) ----
    fo = descriptor_pb2.FileDescriptorProto(name="b200/l2_xpu_infra_mgr.proto", package=OPI, syntax="proto3")
    en = fo.enum_type.add(name="BridgePortType")
    for i, n in enumerate(["BRIDGE_PORT_TYPE_UNSPECIFIED", "BRIDGE_PORT_TYPE_ACCESS", "BRIDGE_PORT_TYPE_TRUNK"]):
        en.value.add(name=n, number=i)
    en = fo.enum_type.add(name="BPOperStatus")
    for i, n in enumerate(["BP_OPER_STATUS_UNSPECIFIED", "BP_OPER_STATUS_UP", "BP_OPER_STATUS_DOWN"]):
        en.value.add(name=n, number=i)
    m = fo.message_type.add(name="BridgePortSpec")
    _field(m, "mac_address", 1, _T.TYPE_BYTES)
    _field(m, "ptype", 2, _T.TYPE_ENUM, type_name=f".{OPI}.BridgePortType")
    _field(m, "logical_bridges", 3, _T.TYPE_STRING, _T.LABEL_REPEATED)
    m = fo.message_type.add(name="BridgePortStatus")
    _field(m, "oper_status", 1, _T.TYPE_ENUM, type_name=f".{OPI}.BPOperStatus")
    m = fo.message_type.add(name="BridgePort")
    _field(m, "name", 1, _T.TYPE_STRING)
    _field(m, "spec", 2, _T.TYPE_MESSAGE, type_name=f".{OPI}.BridgePortSpec")
    _field(m, "status", 3, _T.TYPE_MESSAGE, type_name=f".{OPI}.BridgePortStatus")
    m = fo.message_type.add(name="CreateBridgePortRequest")
    _field(m, "bridge_port_id", 1, _T.TYPE_STRING)
    _field(m, "bridge_port", 2, _T.TYPE_MESSAGE, type_name=f".{OPI}.BridgePort")
    m = fo.message_type.add(name="DeleteBridgePortRequest")
    _field(m, "name", 1, _T.TYPE_STRING)
    _field(m, "allow_missing", 2, _T.TYPE_BOOL)
    pool.Add(fo)
    return pool


_POOL = _build_pool()


def msg(full_name: str):
    """Message class by full name, e.g. ``msg("Vendor.NFRequest")``."""
    return message_factory.GetMessageClass(_POOL.FindMessageTypeByName(full_name))


# (service, method) -> (request type, response type)
METHODS = {
    ("Vendor.LifeCycleService", "Init"): ("Vendor.InitRequest", "Vendor.IpPort"),
    ("Vendor.NetworkFunctionService", "CreateNetworkFunction"): ("Vendor.NFRequest", "Vendor.Empty"),
    ("Vendor.NetworkFunctionService", "DeleteNetworkFunction"): ("Vendor.NFRequest", "Vendor.Empty"),
    ("Vendor.DeviceService", "GetDevices"): ("Vendor.Empty", "Vendor.DeviceListResponse"),
    ("Vendor.DeviceService", "SetNumVfs"): ("Vendor.VfCount", "Vendor.VfCount"),
    (f"{OPI}.BridgePortService", "CreateBridgePort"): (f"{OPI}.CreateBridgePortRequest", f"{OPI}.BridgePort"),
    (f"{OPI}.BridgePortService", "DeleteBridgePort"): (f"{OPI}.DeleteBridgePortRequest", "google.protobuf.Empty"),
}


class B200Vsp:
    """RPC handlers.  State changes are serialised by one lock: the reference VSP assumes serial RPCs
    (it mutates its maps without one, marvell/main.go:361-381); grpc may run handlers concurrently."""

    def __init__(self, device: int = 0, rpm_port: str = "rpm0", nf_devices=None):
        self.device = device
        self.vsp = VspServer(device, rpm_port)
        # stand-in for the veth pairs of ConfigureNetworkInterface (marvell/main.go:105-152,183-215):
        # NF-side MAC -> (NF-side name, dataplane-side name); NoOfPortPairs = 2
        self.nf_devices = nf_devices or {
            "02:00:00:00:01:00": ("nf_interface0", "dp_interface0"),
            "02:00:00:00:01:01": ("nf_interface1", "dp_interface1"),
        }
        self.lock = threading.Lock()
        self.dpu_mode = None
        self.num_vfs = 0

    def _abort(self, ctx, e: Exception):
        ctx.abort(grpc.StatusCode.INTERNAL, str(e))  # Go errors surface as codes.Unknown/Internal with the text

    # ---- LifeCycleService ----
    def Init(self, req, ctx):
        with self.lock:
            try:
                self.dpu_mode = bool(req.dpu_mode)
                if self.dpu_mode:
                    for mac, (_, dp) in self.nf_devices.items():
                        self.vsp.add_nf_device(mac, dp)
                self.vsp.Init(self.dpu_mode)
            except SfcError as e:
                self._abort(ctx, e)
        # doInit returns the address the DPU-side daemon listens on (marvell/main.go:288-292, configureIP)
        return msg("Vendor.IpPort")(ip=IPV6_ADDR_DPU if self.dpu_mode else IPV6_ADDR_HOST, port=DEFAULT_PORT)

    # ---- NetworkFunctionService ----
    def CreateNetworkFunction(self, req, ctx):
        with self.lock:
            try:
                self.vsp.CreateNetworkFunction(req.input, req.output)
            except SfcError as e:
                self._abort(ctx, e)
        return msg("Vendor.Empty")()

    def DeleteNetworkFunction(self, req, ctx):
        with self.lock:
            try:
                self.vsp.DeleteNetworkFunction(req.input, req.output)
            except SfcError as e:
                self._abort(ctx, e)
        return msg("Vendor.Empty")()

    # ---- DeviceService ----
    def GetDevices(self, req, ctx):
        if self.dpu_mode is None:
            ctx.abort(grpc.StatusCode.INTERNAL, "device Store is empty")  # marvell/main.go:625-627
        out = msg("Vendor.DeviceListResponse")()
        if self.dpu_mode:  # the NF-side interfaces (marvell/main.go:628-634)
            # the reference reports the constant "Healthy" (:219-236 derive it from link state); here the string
            # says whether the GPU dataplane behind the interfaces still answers (the device plugin withdraws
            # devices whose health is anything else)
            health = "Healthy" if self.stats() is not None else "Unhealthy"
            for _, (nf_name, _) in sorted(self.nf_devices.items()):
                out.devices[nf_name].ID = nf_name
                out.devices[nf_name].health = health
        else:  # host side: one device per VF (marvell/main.go:635-641); ids are positional here
            for i in range(self.num_vfs):
                vf = f"host-vf{i}"
                out.devices[vf].ID = vf
                out.devices[vf].health = "Healthy"
        return out

    def stats(self):
        """Dataplane hit counters as a dict — packets per action and per stage since the last reset, read back from
        the GPU (SFC_CNT_*; counted by launches that pass SFC_F_COUNTERS).  The p4info of the Intel pipeline declares
        direct counters that p4rt-ctl can read (cmd/intelvsp/p4rt-ctl:1686-1808); the reference daemon never reads
        them, so there is no RPC for this: it is the stat surface of this VSP for whoever embeds it.  Returns None if
        the dataplane does not answer; an offline dataplane (no GPU, rule staging only) reports zeros."""
        from . import consts
        try:
            dp = self.vsp.dataplane
        except SfcError:
            return None
        try:
            c = dp.counters_read(16) if dp.device >= 0 else [0] * 16
        except SfcError:
            return None
        names = ["miss", "acl", "flow", "l2_exact", "l2_wildcard", "lpm4", "lpm6"]
        return dict(to_cp=int(c[consts.CNT_ACTION0]), forward=int(c[consts.CNT_ACTION0 + 1]), drop=int(c[consts.CNT_ACTION0 + 2]),
                    by_stage={n: int(c[consts.CNT_STAGE0 + i]) for i, n in enumerate(names)},
                    table_epoch=dp.table_epoch, table_bytes=dp.table_size)

    def SetNumVfs(self, req, ctx):
        if self.dpu_mode:
            ctx.abort(grpc.StatusCode.INTERNAL, "SetNumVfs is not supported in DPU Mode")  # marvell/main.go:651-653
        if req.vf_cnt < 0:
            ctx.abort(grpc.StatusCode.INTERNAL, "invalid VF Count")
        self.num_vfs = req.vf_cnt
        return msg("Vendor.VfCount")(vf_cnt=req.vf_cnt)

    # ---- BridgePortService ----
    def CreateBridgePort(self, req, ctx):
        with self.lock:
            try:
                name = self.vsp.CreateBridgePort(req.bridge_port.name, bytes(req.bridge_port.spec.mac_address))
            except (SfcError, ValueError) as e:
                self._abort(ctx, e)
        out = msg(f"{OPI}.BridgePort")(name=name)  # marvell/main.go:414-419
        out.spec.CopyFrom(req.bridge_port.spec)
        out.status.SetInParent()
        return out

    def DeleteBridgePort(self, req, ctx):
        with self.lock:
            try:
                self.vsp.DeleteBridgePort(req.name)
            except SfcError as e:
                self._abort(ctx, e)
        return msg("google.protobuf.Empty")()

    # ---- plumbing ----
    def handlers(self):
        by_service = {}
        for (service, method), (req_t, resp_t) in METHODS.items():
            h = grpc.unary_unary_rpc_method_handler(
                getattr(self, method),
                request_deserializer=msg(req_t).FromString,
                response_serializer=lambda m: m.SerializeToString(),
            )
            by_service.setdefault(service, {})[method] = h
        return [grpc.method_handlers_generic_handler(s, hs) for s, hs in by_service.items()]


def serve(address: str = f"unix:{VENDOR_PLUGIN_SOCKET}", device: int = 0, rpm_port: str = "rpm0", max_workers: int = 4):
    """Start the VSP; returns (grpc.Server, B200Vsp).  ``address`` is a grpc target such as
    ``unix:/path/to.sock`` (insecure, like the reference: vendorplugin.go:183-207)."""
    vsp = B200Vsp(device, rpm_port)
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    server.add_generic_rpc_handlers(tuple(vsp.handlers()))
    if server.add_insecure_port(address) == 0 and not address.startswith("unix:"):
        raise RuntimeError(f"could not bind {address}")
    server.start()
    return server, vsp


class NfMacPairer:
    """The caller side of CreateNetworkFunction: how the DPU-side daemon turns CNI ADD / DEL of a network function's
    two interfaces into one RPC each (internal/daemon/dpusidemanager.go:104-139).  MACs are collected per network
    namespace in arrival order; the second ADD calls CreateNetworkFunction(macs[0], macs[1]) — the daemon ignores that
    call's error (:115) — and a DEL that finds two MACs calls DeleteNetworkFunction before dropping the newest one.
    `vsp` is anything with the two methods (a VspClient, a B200Vsp wrapper, a test double)."""

    def __init__(self, vsp):
        self.vsp = vsp
        self.mac_store: dict[str, list[str]] = {}

    def cni_add(self, netns: str, mac: str):
        macs = self.mac_store.setdefault(netns, [])
        macs.append(mac)
        if len(macs) == 2:
            try:
                self.vsp.CreateNetworkFunction(macs[0], macs[1])
            except Exception:  # noqa: BLE001  (dpusidemanager.go:115 drops the error)
                pass

    def cni_del(self, netns: str):
        macs = self.mac_store.get(netns, [])
        if len(macs) == 2:
            self.vsp.DeleteNetworkFunction(macs[0], macs[1])
        self.mac_store[netns] = macs[:-1]


class VspClient:
    """What ``GrpcPlugin`` (internal/daemon/plugin/vendorplugin.go:183-265) does, for tests and demos."""

    def __init__(self, address: str):
        self.channel = grpc.insecure_channel(address)
        self._calls = {}
        for (service, method), (req_t, resp_t) in METHODS.items():
            self._calls[method] = self.channel.unary_unary(
                f"/{service}/{method}",
                request_serializer=lambda m: m.SerializeToString(),
                response_deserializer=msg(resp_t).FromString,
            )

    def __getattr__(self, method):
        try:
            return self.__dict__["_calls"][method]
        except KeyError:
            raise AttributeError(method) from None

    def close(self):
        self.channel.close()


if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser(description="B200 vendor-specific plugin for dpu-operator")
    ap.add_argument("--socket", default=f"unix:{VENDOR_PLUGIN_SOCKET}")
    ap.add_argument("--device", type=int, default=0)
    args = ap.parse_args()
    srv, _ = serve(args.socket, args.device)
    srv.wait_for_termination()
