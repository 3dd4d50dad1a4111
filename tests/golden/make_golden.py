"""Regenerates tests/golden/golden.json from the oracle (run from the repo root:
`python tests/golden/make_golden.py`).  The reference ships no golden vectors for this
path and TFHEpp is unseeded, so the fixture pins the oracle's seeded key
generator and two small gate batches by digest; tests/test_cpu.py checks it on every run."""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402

keys = orc.Keys(20261104)
G = {"seed": 20261104,
     "bk_sha256": hashlib.sha256(keys.bk.tobytes()).hexdigest(),
     "ksk_sha256": hashlib.sha256(keys.ksk.tobytes()).hexdigest()}
for level in (0, 1):
    x0 = orc.encrypt(keys, [0, 1, 0, 1], level, base_index=1)
    x1 = orc.encrypt(keys, [0, 0, 1, 1], level, base_index=101)
    out = orc.gate_batch(keys, orc.OP["NAND"], level, x0, x1)
    G[f"nand_lvl{level}_sha256"] = hashlib.sha256(out.tobytes()).hexdigest()
    G[f"nand_lvl{level}_first_words"] = out[0, :4].tolist()
with open(os.path.join(ROOT, "tests", "golden", "golden.json"), "w") as f:
    json.dump(G, f, indent=1)
print(json.dumps(G, indent=1))
