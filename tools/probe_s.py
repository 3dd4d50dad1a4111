"""A/B of the second-generation blind rotation (k_blind_rotate_s, variants 40..47 = Solinas masks
0, F2, F3, F2|F3, I2|I3, F2|F3|I2|I3, all but I1, all) against the first generation (variant 19):
outputs compared byte for byte with the oracle (small ragged batch incl. extreme ciphertexts, every
gates-per-CTA instantiation) and with variant 19 (1200 random ciphertexts), then CUDA-event timings.
Variant 48 reads key row 0 at every step (timing experiment: key traffic vs arithmetic).
Prints JSON lines.  usage: probe_s.py [count] [variants,comma,separated]"""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import cufhe_b200 as cf  # noqa: E402
from oracle import orc  # noqa: E402  (checker, outside any timed region)

L = cf._lib.lib()
keys = orc.Keys(20261104)
cf.SetGPUNum(1)
cf.Initialize(keys.bk, keys.ksk)
rng = np.random.default_rng(0)
MASKS = {40: 0}
variants = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else sorted(MASKS)


def timed(fn, reps=3, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), sorted(ts)[len(ts) // 2]


def run(variant, g, src, n, dst):
    cf._lib.check(L.b2fhe_set_br_variant(variant))
    cf._lib.check(L.b2fhe_set_gates_per_cta(g))
    cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, dst.data_ptr(), src.data_ptr(), 1 << 29, n))


# ---- parity vs the oracle ----
cnt0 = 13
c0 = orc.encrypt(keys, np.arange(cnt0) & 1, 0)
c0[3, :] = 0
c0[4, :636] = 0xFFFF
c0[5, :] = 0xFFFF
d0 = torch.from_numpy(c0.view(np.int16)).cuda()
refs = np.stack([orc.blind_rotate_extract(keys, c0[i]) for i in range(cnt0)])
d1 = torch.empty((cnt0, 1025), dtype=torch.int32, device="cuda")
ok_all = True
for variant in variants + [5, 0]:
    for g in (0, 1, 2, 3, 4):
        d1.zero_()
        run(variant, g, d0, cnt0, d1)
        torch.cuda.synchronize()
        got = d1.cpu().numpy().view(np.uint32)
        ok = bool(np.array_equal(got, refs))
        ok_all &= ok
        if not ok or g == 0:
            bad = [i for i in range(cnt0) if not np.array_equal(got[i], refs[i])]
            print(json.dumps({"probe": "oracle_parity", "variant": variant, "gates_per_cta": g, "equal": ok, "bad_items": bad}), flush=True)

# ---- byte equality vs the first generation on random ciphertexts ----
cnt = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
x0 = torch.from_numpy(rng.integers(0, 2**16, (cnt, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
nchk = min(cnt, 1200)
ref = torch.zeros((nchk, 1025), dtype=torch.int32, device="cuda")
run(40, 1, x0, nchk, ref)
torch.cuda.synchronize()
for variant in variants:
    got = torch.zeros_like(ref)
    run(variant, 0, x0, nchk, got)
    torch.cuda.synchronize()
    eq = bool(torch.equal(ref, got))
    ok_all &= eq
    print(json.dumps({"probe": "bit_exact_vs_one_gate_per_cta", "variant": variant, "solinas_mask": MASKS[variant], "count": nchk, "equal": eq}), flush=True)

# ---- timings (one launch of gates-per-CTA 4 over the whole batch, no tail split) ----
cf._lib.check(L.b2fhe_set_br_variant(-1))
e = torch.empty((cnt, 1025), dtype=torch.int32, device="cuda")
for variant in variants + [48]:
    best, med = timed(lambda: run(variant, 4, x0, cnt, e))
    print(json.dumps({"probe": "blind_rotate", "variant": variant, "solinas_mask": MASKS.get(variant), "gates_per_cta": 4, "count": cnt,
                      "ms_best": best, "ms_med": med, "gates_per_s": cnt / best * 1e3}), flush=True)
print(json.dumps({"probe": "summary", "all_equal": ok_all}), flush=True)
cf._lib.check(L.b2fhe_set_br_variant(-2))
cf._lib.check(L.b2fhe_set_br_variant(0))
cf._lib.check(L.b2fhe_set_gates_per_cta(0))
cf.CleanUp()
