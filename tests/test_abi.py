"""The C-ABI library loads on a CPU-only box and exports every symbol include/b200sfc.h declares."""
import ctypes
import os
import re

import pytest

from common import ROOT, sfc


def declared_symbols():
    syms = set()
    for name in sorted(os.listdir(os.path.join(ROOT, "include"))):
        text = open(os.path.join(ROOT, "include", name)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        syms |= set(re.findall(r"\b((?:sfc|vsp)_[a-z0-9_]+)\s*\(", text))
    return sorted(syms)


def test_header_declares_something():
    syms = declared_symbols()
    assert len(syms) >= 40 and "sfc_classify" in syms and "sfc_l2_rule_add" in syms


def test_every_declared_symbol_is_exported():
    L = ctypes.CDLL(sfc.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(L, s)]
    assert not missing, f"declared in b200sfc.h but not exported: {missing}"


def test_python_bindings_cover_the_header():
    bound = set(sfc.lib._sfc_signatures)
    assert set(declared_symbols()) <= bound, set(declared_symbols()) - bound


def test_no_cpu_fallback_without_a_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(sfc.SfcError) as e:
        sfc.Dataplane(0)
    assert e.value.code == -19  # -ENODEV


def test_offline_handle_refuses_data_plane_calls():
    dp = sfc.Dataplane(-1)
    dp.commit()
    import numpy as np

    with pytest.raises(sfc.SfcError) as e:
        dp.classify_host(np.zeros((4, 64), np.uint8), np.zeros(4, np.uint16))
    assert e.value.code == -19
    with pytest.raises(sfc.SfcError):
        dp.counters_read()


def test_error_codes_and_messages():
    dp = sfc.Dataplane(-1)
    with pytest.raises(sfc.SfcError) as e:
        dp.port_del("nope")
    assert e.value.code == -2 and "nope" in str(e.value)
    with pytest.raises(sfc.SfcError) as e:
        dp.set_option("flow_load_pct", 1000)
    assert e.value.code == -22
    with pytest.raises(sfc.SfcError) as e:
        dp.lpm_add(4, b"\x0a\x00\x00\x00", 33, 1)
    assert e.value.code == -22


def test_port_registry_semantics():
    # ovs-vsctl --may-exist add-port / del-port / list-ports (ovs-dp/ovsdp.go:40-62,104-112)
    dp = sfc.Dataplane(-1)
    a, b = dp.port_add("eth0"), dp.port_add("vf0-1")
    assert (a, b) == (0, 1)
    assert dp.port_add("eth0") == 0  # --may-exist
    assert dp.port_list() == ["eth0", "vf0-1"]
    dp.port_del("eth0")
    assert dp.port_list() == ["vf0-1"]
    assert dp.port_add("nf-in") == 0  # lowest free id is reused
    assert dp.port_lookup("vf0-1") == 1


def test_header_is_plain_c_and_a_c_program_links_against_the_library(tmp_path):
    """The boundary is a C ABI: a C11 translation unit (what cgo compiles) must take both headers as they
    are, link against the library and drive the control plane of an offline handle."""
    import shutil
    import subprocess
    if not shutil.which("gcc"):
        pytest.skip("no gcc")
    src = tmp_path / "abi.c"
    src.write_text(r'''
#include <stdio.h>
#include <string.h>
#include "b200sfc.h"
#include "b200vsp.h"
int main(void) {
    sfc_handle* h = NULL;
    if (sfc_create(-1, &h) != 0) return 1;
    uint16_t a = 0, b = 0;
    if (sfc_port_add(h, "vf0", &a) || sfc_port_add(h, "nf0-in", &b)) return 2;
    if (sfc_l2_rule_add(h, a, b, NULL, NULL)) return 3;
    sfc_batch batch; memset(&batch, 0, sizeof batch);
    if (sizeof(sfc_batch) != 32 || sizeof(sfc_flow_key) != 40 || sizeof(sfc_acl_rule) != 88) return 4;
    if (sfc_commit(h)) return 5;
    if (sfc_classify(h, &batch, &a, (uint64_t*)&batch, 1, 0, NULL) == 0) return 6;  /* offline: must refuse */
    printf("%s|%zu|%s\n", sfc_version(), sfc_table_size(h), sfc_last_error(h));
    return sfc_destroy(h);
}
''')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.join(root, "dpu-operator_b200")
    exe = tmp_path / "abi"
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(root, "include"), str(src), "-o", str(exe),
                    "-L", libdir, "-lb200sfc", f"-Wl,-rpath,{libdir}"], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.strip().split("|")
    assert out[0].startswith("b200sfc") and int(out[1]) > 512 and "offline" in out[2]


def test_a_deleted_port_takes_its_steering_rules_with_it():
    """Port ids are recycled (lowest free id first).  A port added after a delete must not inherit the old
    port's rules: sfc_port_del drops the L2 rules naming the id (OVS: flows naming a removed port stop matching)."""
    import numpy as np
    from common import Scenario, sfc
    from test_image_walk import Image

    sc = Scenario(sfc.Dataplane(-1))
    dp = sc.dp
    a, b, c = dp.port_add("a"), dp.port_add("b"), dp.port_add("c")
    mac = bytes.fromhex("020000000001")
    dp.l2_rule_add(b, c)          # in_port = b
    dp.l2_rule_add(a, b, mac)     # out_port = b
    dp.l2_rule_add(a, c)          # does not name b
    dp.commit()
    img = Image(dp.table_export())
    assert img.l2(b, bytes(6)) is not None and (img.l2(a, mac) >> 32) & 0xFFFF == b
    dp.port_del("b")
    assert dp.port_add("new") == b  # the id is recycled
    dp.commit()
    img = Image(dp.table_export())
    assert img.l2(b, bytes(6)) is None, "the new port inherited the deleted port's wildcard rule"
    assert (img.l2(a, mac) >> 32) & 0xFFFF == c and (img.l2(a, mac) >> 28) & 0xF == sfc.consts.STAGE_L2_WILD  # the MAC rule towards b is gone
    assert dp.stat("full_commits") == 1  # L2 edits go into the idle copy, no rebuild
