"""Regenerates tests/golden/wire_golden.json from the REAL TFHEpp (oracle/_ref, needs
/root/reference): the framing bytes of an EvalKey archive written by TFHEpp's own cereal
serialisation -- everything except the two big key arrays -- and one complete 2-element
ciphertext-vector archive per level.  tests/test_cpu.py checks cufhe_b200's writer and reader
against these on machines without the reference.  Run from the repo root:
`python tests/golden/make_wire_golden.py`."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import tfhepp  # noqa: E402

BK, KSK = 636 * 4 * 2 * 1024 * 4, 1024 * 7 * 2 * 637 * 2
tfhepp.Keys()
ek = tfhepp.serialize_evalkey(0)
head = ek[:117].copy()                       # flag, lweParams, id of tuple member 0
head[33:41] = 0                              # lvl1.alpha: uninitialised in TFHEpp (lweParams.hpp:43-50)
mid = ek[117 + BK:117 + BK + 48]             # members 1..11 (null) and the id of member 12
tail = ek[117 + BK + 48 + KSK:]              # members 13..27
G = {"evalkey_bytes": int(ek.size), "evalkey_head_hex": head.tobytes().hex(), "evalkey_mid_hex": mid.tobytes().hex(),
     "evalkey_tail_hex": tail.tobytes().hex()}
for level in (0, 1):
    cts = tfhepp.encrypt([1, 0], level)
    G[f"tlwes_lvl{level}_hex"] = tfhepp.serialize_tlwes(cts, level).tobytes().hex()
with open(os.path.join(ROOT, "tests", "golden", "wire_golden.json"), "w") as f:
    json.dump(G, f, indent=1)
print({k: (v if isinstance(v, int) else len(v) // 2) for k, v in G.items()})
