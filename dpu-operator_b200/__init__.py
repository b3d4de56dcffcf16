"""dpu-operator_b200 — B200-native SFC flow classification behind the dpu-operator plugin API.

The product is the C-ABI library ``libb200sfc.so`` (declared in ``include/b200sfc.h``, built from
``csrc/``).  This package is the thin Python view of that ABI used by the tests, ``bench.py`` and
the VSP front door.  There is no CPU fallback anywhere in it: if the library has not been built
the import fails, and if no B200 is present ``Dataplane()`` raises.

The directory name carries a hyphen (it is the repo's name), so import it with
``importlib.import_module("dpu-operator_b200")``.
"""
from __future__ import annotations

from .capi import (  # noqa: F401
    ACL_RULE_DTYPE,
    BATCH_DTYPE,
    FLOW_ACTION_DTYPE,
    FLOW_KEY_DTYPE,
    GEN_SPEC_DTYPE,
    LIB_PATH,
    Dataplane,
    Group,
    SfcError,
    STATIC_AREA_BYTES,
    VspServer,
    classify_static,
    gen_acl_rules,
    gen_flows,
    gen_mac,
    gen_packets_host,
    gen_prefixes,
    lib,
    make_gen_spec,
    result_fields,
)
from . import consts  # noqa: F401

__all__ = [
    "Dataplane",
    "Group",
    "SfcError",
    "lib",
    "LIB_PATH",
    "consts",
    "FLOW_KEY_DTYPE",
    "FLOW_ACTION_DTYPE",
    "ACL_RULE_DTYPE",
    "BATCH_DTYPE",
    "GEN_SPEC_DTYPE",
    "make_gen_spec",
    "gen_flows",
    "gen_packets_host",
    "gen_prefixes",
    "gen_acl_rules",
    "gen_mac",
    "result_fields",
    "classify_static",
    "VspServer",
    "STATIC_AREA_BYTES",
]
