"""Regenerate tests/golden/handmade_cases.json.

The inputs are the hand-built frames of tests/packets.py; the expected result words come from the
independent Python oracle (oracle/py_oracle.py), and each case gets a short note decoded from the
result so a reviewer can check it against the pinned semantics by eye.  Run from the repo root:
    python tests/golden/make_golden.py
(no GPU and no /root/reference needed; the reference holds no per-packet vectors to import —
see SURVEY.md §4: "Not one test asserts a rule string, a table entry, or a per-packet decision").
"""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from common import Scenario, sfc  # noqa: E402
from packets import handmade_cases  # noqa: E402

STAGES = ["miss", "acl", "flow", "l2-exact", "l2-wildcard", "lpm4", "lpm6"]
ACTIONS = ["to_cp", "forward", "drop"]

sc = Scenario(sfc.Dataplane(-1), with_py=True)
hdr, ports = handmade_cases(sc)
cases = []
for i in range(len(ports)):
    r = sc.py.classify(bytes(hdr[i]), int(ports[i]))
    a, nf, p, st, idx = (int(x[0]) for x in sfc.result_fields([r]))
    note = f"{ACTIONS[a]} via {STAGES[st]}#{idx} -> port {p}, next_nf {nf if nf != 255 else 'none'}"
    cases.append(dict(hdr=bytes(hdr[i]).hex(), in_port=int(ports[i]), result=f"{r:#018x}", note=note))
out = os.path.join(os.path.dirname(__file__), "handmade_cases.json")
json.dump(dict(source="tests/packets.py + oracle/py_oracle.py", cases=cases), open(out, "w"), indent=1)
print(f"wrote {len(cases)} cases to {out}")
