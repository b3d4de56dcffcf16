"""ctypes view of libb200sfc.so (include/b200sfc.h).

Every method of ``Dataplane`` is a one-to-one call of the C entry point of the same name; the
argument meaning and the error behaviour (negative errno + ``sfc_last_error``) are the library's.
Device buffers are passed as integer addresses (e.g. ``torch.Tensor.data_ptr()``): torch is the
plumbing that owns device memory and streams in the tests and the bench, not part of the ABI.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import consts

LIB_PATH = os.environ.get("B200SFC_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb200sfc.so")

FLOW_KEY_DTYPE = np.dtype(
    [("src", "u1", 16), ("dst", "u1", 16), ("sport", "<u2"), ("dport", "<u2"), ("vlan", "<u2"), ("proto", "u1"),
     ("ip_ver", "u1")]
)
FLOW_ACTION_DTYPE = np.dtype([("action", "u1"), ("reserved", "u1"), ("out_port", "<u2")])
ACL_RULE_DTYPE = np.dtype(
    [("src", "u1", 16), ("src_mask", "u1", 16), ("dst", "u1", 16), ("dst_mask", "u1", 16),
     ("in_port", "<u2"), ("in_port_mask", "<u2"), ("vlan", "<u2"), ("vlan_mask", "<u2"),
     ("sport", "<u2"), ("sport_mask", "<u2"), ("dport", "<u2"), ("dport_mask", "<u2"),
     ("proto", "u1"), ("proto_mask", "u1"), ("ip_ver", "u1"), ("verdict", "u1"), ("priority", "<u4")]
)
BATCH_DTYPE = np.dtype([("hdr", "<u8"), ("in_port", "<u8"), ("result", "<u8"), ("n", "<u4"), ("reserved", "<u4")])
GEN_SPEC_DTYPE = np.dtype(
    [("seed_pkts", "<u8"), ("seed_rules", "<u8"), ("seed_lpm", "<u8"), ("seed_acl", "<u8"),
     ("n_flows", "<u4"), ("n_ports", "<u4"), ("port_base", "<u4"), ("n_out_ports", "<u4"),
     ("hit_permille", "<u4"), ("nonip_permille", "<u4"), ("v6_permille", "<u4"), ("vlan_permille", "<u4"),
     ("zipf", "<u4"), ("n_macs", "<u4"), ("n_pfx4", "<u4"), ("n_pfx6", "<u4"), ("lpm_permille", "<u4"),
     ("n_acl", "<u4")]
)
assert FLOW_KEY_DTYPE.itemsize == 40 and FLOW_ACTION_DTYPE.itemsize == 4
assert ACL_RULE_DTYPE.itemsize == 88 and GEN_SPEC_DTYPE.itemsize == 88


class SfcError(RuntimeError):
    """A C-ABI call returned a negative errno; ``.code`` holds it."""

    def __init__(self, code: int, msg: str):
        super().__init__(f"[{code}] {msg}")
        self.code = code


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C dpu-operator_b200/csrc`. There is no CPU fallback."
        )
    L = C.CDLL(LIB_PATH)
    vp, u16p, u32p, u64p, u8p = C.c_void_p, C.POINTER(C.c_uint16), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.c_void_p
    sig = {
        "sfc_create": (C.c_int, [C.c_int, C.POINTER(vp)]),
        "sfc_destroy": (C.c_int, [vp]),
        "sfc_last_error": (C.c_char_p, [vp]),
        "sfc_version": (C.c_char_p, []),
        "sfc_set_option": (C.c_int, [vp, C.c_char_p, C.c_long]),
        "sfc_port_add": (C.c_int, [vp, C.c_char_p, u16p]),
        "sfc_port_del": (C.c_int, [vp, C.c_char_p]),
        "sfc_port_lookup": (C.c_int, [vp, C.c_char_p, u16p]),
        "sfc_port_list": (C.c_long, [vp, C.c_char_p, C.c_size_t]),
        "sfc_l2_rule_add": (C.c_int, [vp, C.c_uint16, C.c_uint16, u8p, u32p]),
        "sfc_l2_rule_del": (C.c_int, [vp, C.c_uint16, u8p]),
        "sfc_chain_set": (C.c_int, [vp, vp, vp, C.c_size_t]),
        "sfc_flow_add": (C.c_int, [vp, vp, vp, u32p]),
        "sfc_flow_add_bulk": (C.c_int, [vp, vp, vp, C.c_size_t, u32p]),
        "sfc_flow_del": (C.c_int, [vp, vp]),
        "sfc_lpm_add": (C.c_int, [vp, C.c_int, u8p, C.c_int, C.c_uint16, u32p]),
        "sfc_lpm_del": (C.c_int, [vp, C.c_int, u8p, C.c_int]),
        "sfc_acl_add": (C.c_int, [vp, vp, u32p]),
        "sfc_acl_del": (C.c_int, [vp, C.c_uint32]),
        "sfc_clear": (C.c_int, [vp]),
        "sfc_commit": (C.c_int, [vp]),
        "sfc_commit_stream": (C.c_int, [vp, vp]),
        "sfc_stat": (C.c_long, [vp, C.c_char_p]),
        "sfc_table_size": (C.c_size_t, [vp]),
        "sfc_table_epoch": (C.c_uint64, [vp]),
        "sfc_table_export": (C.c_int, [vp, vp, C.c_size_t]),
        "sfc_table_import": (C.c_int, [vp, vp, C.c_size_t]),
        "sfc_table_export_device": (C.c_int, [vp, vp, C.c_size_t, vp]),
        "sfc_table_import_device": (C.c_int, [vp, vp, C.c_size_t, vp]),
        "sfc_classify": (C.c_int, [vp, vp, vp, vp, C.c_size_t, C.c_uint32, vp]),
        "sfc_classify_batches": (C.c_int, [vp, vp, C.c_uint32, C.c_uint32, vp]),
        "sfc_classify_host": (C.c_int, [vp, vp, vp, vp, C.c_size_t, C.c_uint32]),
        "sfc_classify_host_submit": (C.c_int, [vp, vp, vp, vp, C.c_size_t, C.c_uint32, u64p]),
        "sfc_classify_host_wait": (C.c_int, [vp, C.c_uint64]),
        "sfc_sync": (C.c_int, [vp, vp]),
        "sfc_group_create": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(vp)]),
        "sfc_group_destroy": (C.c_int, [vp]),
        "sfc_group_size": (C.c_int, [vp]),
        "sfc_group_handle": (vp, [vp, C.c_int]),
        "sfc_group_broadcast_table": (C.c_int, [vp]),
        "sfc_group_reduce_counters": (C.c_int, [vp]),
        "sfc_group_last_ms": (C.c_float, [vp, C.c_int]),
        "sfc_group_last_error": (C.c_char_p, [vp]),
        "sfc_counters_read": (C.c_int, [vp, u64p, C.c_size_t]),
        "sfc_counters_reset": (C.c_int, [vp]),
        "sfc_counters_device_ptr": (vp, [vp]),
        "sfc_launch_count": (C.c_uint64, [vp]),
        "sfc_last_kernel_ms": (C.c_float, [vp]),
        "sfc_classify_static": (C.c_int, [C.c_int, vp, C.c_size_t, vp, vp, vp, C.c_size_t, vp, C.c_char_p, C.c_size_t]),
        "sfc_gen_packets": (C.c_int, [vp, vp, C.c_uint64, C.c_size_t, vp, vp, vp]),
        "sfc_gen_packets_host": (None, [vp, C.c_uint64, C.c_size_t, vp, vp]),
        "sfc_gen_flows": (None, [vp, C.c_uint32, C.c_size_t, vp, vp]),
        "sfc_gen_mac": (None, [C.c_uint32, vp]),
        "sfc_gen_prefix": (None, [vp, C.c_int, C.c_uint32, vp, C.POINTER(C.c_int), u16p]),
        "sfc_gen_acl_rule": (None, [vp, C.c_uint32, vp]),
        # include/b200vsp.h
        "vsp_create": (C.c_int, [C.c_int, C.c_char_p, C.POINTER(vp)]),
        "vsp_destroy": (C.c_int, [vp]),
        "vsp_last_error": (C.c_char_p, [vp]),
        "vsp_add_nf_device": (C.c_int, [vp, C.c_char_p, C.c_char_p]),
        "vsp_init": (C.c_int, [vp, C.c_int]),
        "vsp_create_bridge_port": (C.c_int, [vp, C.c_char_p, vp, C.c_char_p, C.c_size_t]),
        "vsp_delete_bridge_port": (C.c_int, [vp, C.c_char_p]),
        "vsp_create_network_function": (C.c_int, [vp, C.c_char_p, C.c_char_p]),
        "vsp_delete_network_function": (C.c_int, [vp, C.c_char_p, C.c_char_p]),
        "vsp_flow_dump": (C.c_long, [vp, C.c_char_p, C.c_size_t]),
        "vsp_dataplane": (vp, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)  # AttributeError here = the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    L._sfc_signatures = sig
    return L


lib = _load()


def _ptr(a: np.ndarray) -> int:
    return a.ctypes.data


def _mac_buf(mac):
    if mac is None:
        return None
    if isinstance(mac, str):
        mac = bytes(int(x, 16) for x in mac.split(":"))
    mac = bytes(mac)
    if len(mac) != 6:
        raise ValueError("MAC must be 6 bytes")
    return C.cast(C.create_string_buffer(mac, 6), C.c_void_p)


class Dataplane:
    """One ``sfc_handle``.  Method names are the C entry points without the ``sfc_`` prefix."""

    def __init__(self, device: int = 0, _borrowed=None):
        self._owned = _borrowed is None
        self.device = device
        if _borrowed is not None:
            self._h = C.c_void_p(_borrowed)
            return
        h = C.c_void_p()
        rc = lib.sfc_create(device, C.byref(h))
        if rc:
            raise SfcError(rc, "sfc_create failed (no CUDA device of compute capability 10.x? there is no CPU fallback)")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            if self._owned:
                lib.sfc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc: int):
        if rc < 0:
            raise SfcError(rc, lib.sfc_last_error(self._h).decode())
        return rc

    # ---- options / ports ----
    def set_option(self, name: str, value: int):
        self._chk(lib.sfc_set_option(self._h, name.encode(), value))

    def port_add(self, name: str) -> int:
        pid = C.c_uint16()
        self._chk(lib.sfc_port_add(self._h, name.encode(), C.byref(pid)))
        return pid.value

    def port_del(self, name: str):
        self._chk(lib.sfc_port_del(self._h, name.encode()))

    def port_lookup(self, name: str) -> int:
        pid = C.c_uint16()
        rc = lib.sfc_port_lookup(self._h, name.encode(), C.byref(pid))
        if rc:
            raise SfcError(rc, f"no port named {name!r}")
        return pid.value

    def port_list(self) -> list[str]:
        need = lib.sfc_port_list(self._h, None, 0)
        buf = C.create_string_buffer(need)
        lib.sfc_port_list(self._h, buf, need)
        return [s for s in buf.value.decode().split("\n") if s]

    # ---- L2 steering / chain ----
    def l2_rule_add(self, in_port: int, out_port: int, dst_mac=None) -> int:
        rid = C.c_uint32()
        self._chk(lib.sfc_l2_rule_add(self._h, in_port, out_port, _mac_buf(dst_mac), C.byref(rid)))
        return rid.value

    def l2_rule_del(self, in_port: int, dst_mac=None):
        self._chk(lib.sfc_l2_rule_del(self._h, in_port, _mac_buf(dst_mac)))

    def chain_set(self, nf_in_ports, nf_out_ports):
        a = np.ascontiguousarray(nf_in_ports, dtype=np.uint16)
        b = np.ascontiguousarray(nf_out_ports, dtype=np.uint16)
        assert a.shape == b.shape
        self._chk(lib.sfc_chain_set(self._h, _ptr(a), _ptr(b), a.size))

    # ---- flows ----
    def flow_add_bulk(self, keys: np.ndarray, acts: np.ndarray) -> int:
        keys = np.ascontiguousarray(keys, dtype=FLOW_KEY_DTYPE)
        acts = np.ascontiguousarray(acts, dtype=FLOW_ACTION_DTYPE)
        assert keys.shape == acts.shape
        first = C.c_uint32()
        self._chk(lib.sfc_flow_add_bulk(self._h, _ptr(keys), _ptr(acts), keys.size, C.byref(first)))
        return first.value

    def flow_add(self, key, act) -> int:
        keys = np.ascontiguousarray(key, dtype=FLOW_KEY_DTYPE).reshape(1)
        acts = np.ascontiguousarray(act, dtype=FLOW_ACTION_DTYPE).reshape(1)
        fid = C.c_uint32()
        self._chk(lib.sfc_flow_add(self._h, _ptr(keys), _ptr(acts), C.byref(fid)))
        return fid.value

    def flow_del(self, key):
        keys = np.ascontiguousarray(key, dtype=FLOW_KEY_DTYPE).reshape(1)
        self._chk(lib.sfc_flow_del(self._h, _ptr(keys)))

    # ---- LPM / ACL ----
    def lpm_add(self, af: int, prefix: bytes, length: int, out_port: int) -> int:
        pid = C.c_uint32()
        buf = C.create_string_buffer(bytes(prefix).ljust(16, b"\0"), 16)
        self._chk(lib.sfc_lpm_add(self._h, af, C.cast(buf, C.c_void_p), length, out_port, C.byref(pid)))
        return pid.value

    def lpm_del(self, af: int, prefix: bytes, length: int):
        buf = C.create_string_buffer(bytes(prefix).ljust(16, b"\0"), 16)
        self._chk(lib.sfc_lpm_del(self._h, af, C.cast(buf, C.c_void_p), length))

    def acl_add(self, rule) -> int:
        r = np.ascontiguousarray(rule, dtype=ACL_RULE_DTYPE).reshape(1)
        rid = C.c_uint32()
        self._chk(lib.sfc_acl_add(self._h, _ptr(r), C.byref(rid)))
        return rid.value

    def acl_del(self, rule_id: int):
        self._chk(lib.sfc_acl_del(self._h, rule_id))

    def clear(self):
        self._chk(lib.sfc_clear(self._h))

    # ---- commit / image ----
    def commit(self):
        self._chk(lib.sfc_commit(self._h))

    def commit_stream(self, stream: int = 0):
        self._chk(lib.sfc_commit_stream(self._h, stream))

    def stat(self, name: str) -> int:
        v = lib.sfc_stat(self._h, name.encode())
        if v < 0:
            raise SfcError(int(v), f"unknown stat {name!r}")
        return int(v)

    @property
    def table_size(self) -> int:
        return lib.sfc_table_size(self._h)

    @property
    def table_epoch(self) -> int:
        return lib.sfc_table_epoch(self._h)

    def table_export(self) -> np.ndarray:
        buf = np.empty(self.table_size, dtype=np.uint8)
        self._chk(lib.sfc_table_export(self._h, _ptr(buf), buf.size))
        return buf

    def table_import(self, image: np.ndarray):
        image = np.ascontiguousarray(image, dtype=np.uint8)
        self._chk(lib.sfc_table_import(self._h, _ptr(image), image.size))

    def table_export_device(self, dptr: int, nbytes: int, stream: int = 0):
        self._chk(lib.sfc_table_export_device(self._h, dptr, nbytes, stream))

    def table_import_device(self, dptr: int, nbytes: int, stream: int = 0):
        self._chk(lib.sfc_table_import_device(self._h, dptr, nbytes, stream))

    # ---- data plane ----
    def classify(self, hdr_ptr: int, in_port_ptr: int, result_ptr: int, n: int, flags: int = 0, stream: int = 0):
        self._chk(lib.sfc_classify(self._h, hdr_ptr, in_port_ptr, result_ptr, n, flags, stream))

    def classify_batches(self, batches, flags: int = 0, stream: int = 0):
        """``sfc_classify_batches``: ``batches`` is a sequence of (hdr_ptr, in_port_ptr, result_ptr, n)
        device addresses, or a ready BATCH_DTYPE array."""
        if not (isinstance(batches, np.ndarray) and batches.dtype == BATCH_DTYPE):
            arr = np.zeros(len(batches), dtype=BATCH_DTYPE)
            for i, (hp, pp, rp, n) in enumerate(batches):
                arr[i] = (hp, pp, rp, n, 0)
            batches = arr
        self._chk(lib.sfc_classify_batches(self._h, _ptr(batches), batches.size, flags, stream))

    def classify_host(self, hdr: np.ndarray, in_port: np.ndarray, flags: int = 0, out: np.ndarray | None = None):
        hdr = np.ascontiguousarray(hdr, dtype=np.uint8)
        in_port = np.ascontiguousarray(in_port, dtype=np.uint16)
        n = in_port.size
        assert hdr.size == n * consts.HDR_BYTES
        if out is None:
            out = np.empty(n, dtype=np.uint64)
        self._chk(lib.sfc_classify_host(self._h, _ptr(hdr), _ptr(in_port), _ptr(out), n, flags))
        return out

    def classify_host_ptr(self, hdr_ptr: int, in_port_ptr: int, result_ptr: int, n: int, flags: int = 0):
        """Same call with raw host addresses (e.g. pinned torch tensors)."""
        self._chk(lib.sfc_classify_host(self._h, hdr_ptr, in_port_ptr, result_ptr, n, flags))

    def classify_host_submit(self, hdr_ptr: int, in_port_ptr: int, result_ptr: int, n: int, flags: int = 0) -> int:
        """Queue one batch held in host memory (raw addresses); returns the ticket for classify_host_wait."""
        t = C.c_uint64(0)
        self._chk(lib.sfc_classify_host_submit(self._h, hdr_ptr, in_port_ptr, result_ptr, n, flags, C.byref(t)))
        return t.value

    def classify_host_wait(self, ticket: int = 0xFFFFFFFFFFFFFFFF):
        self._chk(lib.sfc_classify_host_wait(self._h, ticket))

    def sync(self, stream: int = 0):
        self._chk(lib.sfc_sync(self._h, stream))

    def counters_read(self, n: int = consts.CNT_TOTAL) -> np.ndarray:
        out = np.zeros(n, dtype=np.uint64)
        self._chk(lib.sfc_counters_read(self._h, out.ctypes.data_as(C.POINTER(C.c_uint64)), n))
        return out

    def counters_reset(self):
        self._chk(lib.sfc_counters_reset(self._h))

    @property
    def counters_device_ptr(self) -> int:
        return lib.sfc_counters_device_ptr(self._h)

    @property
    def launch_count(self) -> int:
        return lib.sfc_launch_count(self._h)

    @property
    def last_kernel_ms(self) -> float:
        return lib.sfc_last_kernel_ms(self._h)

    def gen_packets(self, spec: np.ndarray, first: int, n: int, hdr_ptr: int, in_port_ptr: int, stream: int = 0):
        self._chk(lib.sfc_gen_packets(self._h, _ptr(spec), first, n, hdr_ptr, in_port_ptr, stream))


class VspServer:
    """The host-side mirror of the reference's Marvell vendor plugin (include/b200vsp.h).  Method
    names are the RPCs of dpu-api/api.proto and OPI BridgePortService."""

    def __init__(self, device: int = 0, rpm_port: str = "rpm0"):
        v = C.c_void_p()
        rc = lib.vsp_create(device, rpm_port.encode(), C.byref(v))
        if rc:
            raise SfcError(rc, "vsp_create failed")
        self._v, self.device = v, device

    def __del__(self):
        if getattr(self, "_v", None):
            lib.vsp_destroy(self._v)
            self._v = None

    def _chk(self, rc):
        if rc < 0:
            raise SfcError(rc, lib.vsp_last_error(self._v).decode())

    def add_nf_device(self, nf_mac: str, dp_interface: str):
        self._chk(lib.vsp_add_nf_device(self._v, nf_mac.encode(), dp_interface.encode()))

    def Init(self, dpu_mode: bool = True):
        self._chk(lib.vsp_init(self._v, int(dpu_mode)))

    def CreateBridgePort(self, name: str, mac: bytes) -> str:
        out = C.create_string_buffer(128)
        self._chk(lib.vsp_create_bridge_port(self._v, name.encode(), _mac_buf(mac), out, len(out)))
        return out.value.decode()

    def DeleteBridgePort(self, name: str):
        self._chk(lib.vsp_delete_bridge_port(self._v, name.encode()))

    def CreateNetworkFunction(self, input_mac: str, output_mac: str):
        self._chk(lib.vsp_create_network_function(self._v, input_mac.encode(), output_mac.encode()))

    def DeleteNetworkFunction(self, input_mac: str, output_mac: str):
        self._chk(lib.vsp_delete_network_function(self._v, input_mac.encode(), output_mac.encode()))

    def flow_dump(self) -> list[str]:
        need = lib.vsp_flow_dump(self._v, None, 0)
        buf = C.create_string_buffer(need)
        lib.vsp_flow_dump(self._v, buf, need)
        return [x for x in buf.value.decode().split("\n") if x]

    @property
    def dataplane(self) -> "Dataplane":
        h = lib.vsp_dataplane(self._v)
        if not h:
            raise SfcError(-19, "VSP has no dataplane yet (call Init(dpu_mode=True))")
        return Dataplane(self.device, _borrowed=h)


STATIC_AREA_BYTES = 4 << 20


class Group:
    """``sfc_group``: one process driving several GPUs.  ``group[r]`` is the Dataplane of rank r (borrowed: the
    group owns the handles); rules are staged and committed on ``group[0]`` and replicated with ``broadcast_table``."""

    def __init__(self, n_gpus: int, devices=None):
        arr = (C.c_int * n_gpus)(*devices) if devices is not None else None
        g = C.c_void_p()
        rc = lib.sfc_group_create(n_gpus, arr, C.byref(g))
        if rc:
            raise SfcError(rc, "sfc_group_create failed (fewer GPUs than asked for, or NCCL cannot be opened)")
        self._g = g
        self.ranks = [Dataplane(devices[r] if devices is not None else r, _borrowed=lib.sfc_group_handle(g, r)) for r in range(n_gpus)]

    def __len__(self):
        return len(self.ranks)

    def __getitem__(self, r):
        return self.ranks[r]

    def _chk(self, rc):
        if rc < 0:
            raise SfcError(rc, lib.sfc_group_last_error(self._g).decode())

    def broadcast_table(self) -> float:
        self._chk(lib.sfc_group_broadcast_table(self._g))
        return lib.sfc_group_last_ms(self._g, 0)

    def reduce_counters(self) -> float:
        self._chk(lib.sfc_group_reduce_counters(self._g))
        return lib.sfc_group_last_ms(self._g, 1)

    def close(self):
        if getattr(self, "_g", None):
            for d in self.ranks:
                d._h = None
            lib.sfc_group_destroy(self._g)
            self._g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def classify_static(device: int, image, hdr_ptr: int, in_port_ptr: int, result_ptr: int, n: int, stream: int = 0):
    """``sfc_classify_static``: classify with no handle and no allocation.  ``image`` is a host
    uint8 array holding a table image, or None to reuse the one the previous call left on the
    device.  Raises SfcError with the library's message on failure."""
    err = C.create_string_buffer(512)
    if image is None:
        ptr, nbytes = None, 0
    else:
        image = np.ascontiguousarray(image, dtype=np.uint8)
        ptr, nbytes = image.ctypes.data, image.size
    rc = lib.sfc_classify_static(device, ptr, nbytes, hdr_ptr, in_port_ptr, result_ptr, n, stream, err, len(err))
    if rc:
        raise SfcError(rc, err.value.decode())


# ---------------------------------------------------------------- synthetic workload (host) ----
def make_gen_spec(**kw) -> np.ndarray:
    """An ``sfc_gen_spec`` with the SURVEY.md §8(d) seeds; keyword arguments override fields."""
    s = np.zeros(1, dtype=GEN_SPEC_DTYPE)
    s["seed_pkts"], s["seed_rules"] = consts.SEED_PKTS, consts.SEED_RULES
    s["seed_lpm"], s["seed_acl"] = consts.SEED_LPM, consts.SEED_ACL
    s["n_ports"], s["n_out_ports"], s["n_macs"] = 8, 8, 8
    for k, v in kw.items():
        s[k] = v
    return s


def gen_flows(spec: np.ndarray, first: int, n: int):
    keys = np.zeros(n, dtype=FLOW_KEY_DTYPE)
    acts = np.zeros(n, dtype=FLOW_ACTION_DTYPE)
    lib.sfc_gen_flows(_ptr(spec), first, n, _ptr(keys), _ptr(acts))
    return keys, acts


def gen_packets_host(spec: np.ndarray, first: int, n: int):
    hdr = np.zeros((n, consts.HDR_BYTES), dtype=np.uint8)
    ports = np.zeros(n, dtype=np.uint16)
    lib.sfc_gen_packets_host(_ptr(spec), first, n, _ptr(hdr), _ptr(ports))
    return hdr, ports


def gen_mac(i: int) -> bytes:
    buf = C.create_string_buffer(6)
    lib.sfc_gen_mac(i, C.cast(buf, C.c_void_p))
    return buf.raw


def gen_prefixes(spec: np.ndarray, af: int, n: int):
    """Unique prefixes of family af among the first n draws: list of (bytes, len, out_port)."""
    out, seen = [], set()
    buf = C.create_string_buffer(16)
    ln, port = C.c_int(), C.c_uint16()
    for j in range(n):
        lib.sfc_gen_prefix(_ptr(spec), af, j, C.cast(buf, C.c_void_p), C.byref(ln), C.byref(port))
        p = buf.raw[: 4 if af == 4 else 16]
        if (p, ln.value) in seen:
            continue
        seen.add((p, ln.value))
        out.append((p, ln.value, port.value))
    return out


def gen_acl_rules(spec: np.ndarray, n: int) -> np.ndarray:
    rules = np.zeros(n, dtype=ACL_RULE_DTYPE)
    for j in range(n):
        lib.sfc_gen_acl_rule(_ptr(spec), j, rules[j : j + 1].ctypes.data)
    return rules


def result_fields(r):
    """Split result words into (action, next_nf, out_port, stage, index) arrays."""
    r = np.asarray(r, dtype=np.uint64)
    return (
        (r >> np.uint64(56)).astype(np.uint8),
        ((r >> np.uint64(48)) & np.uint64(0xFF)).astype(np.uint8),
        ((r >> np.uint64(32)) & np.uint64(0xFFFF)).astype(np.uint16),
        ((r >> np.uint64(28)) & np.uint64(0xF)).astype(np.uint8),
        (r & np.uint64(0x0FFFFFFF)).astype(np.uint32),
    )
