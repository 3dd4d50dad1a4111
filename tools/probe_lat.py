"""Latency kernel (variant 5 = one gate per 4-CTA cluster) against
the oracle (13 ciphertexts incl. extreme ones) and against the throughput kernel (variant 40, one gate per
CTA) on random ciphertexts, then CUDA-event timings of one blind-rotation launch at small batch sizes and
of a whole NAND gate (blind rotation + key switch) through b2fhe_gate_batch in latency mode.
Prints JSON lines.  usage: probe_lat.py"""
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


def timed(fn, reps=7, warm=2):
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


cnt0 = 13
c0 = orc.encrypt(keys, np.arange(cnt0) & 1, 0)
c0[3, :] = 0
c0[4, :636] = 0xFFFF
c0[5, :] = 0xFFFF
d0 = torch.from_numpy(c0.view(np.int16)).cuda()
refs = np.stack([orc.blind_rotate_extract(keys, c0[i]) for i in range(cnt0)])
d1 = torch.empty((cnt0, 1025), dtype=torch.int32, device="cuda")
ok_all = True
for variant in (5,):
    d1.zero_()
    run(variant, 0, d0, cnt0, d1)
    torch.cuda.synchronize()
    got = d1.cpu().numpy().view(np.uint32)
    ok = bool(np.array_equal(got, refs))
    ok_all &= ok
    print(json.dumps({"probe": "oracle_parity", "variant": variant, "equal": ok,
                      "bad_items": [i for i in range(cnt0) if not np.array_equal(got[i], refs[i])]}), flush=True)

cnt = 296
x0 = torch.from_numpy(rng.integers(0, 2**16, (cnt, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
ref = torch.zeros((cnt, 1025), dtype=torch.int32, device="cuda")
run(40, 1, x0, cnt, ref)
torch.cuda.synchronize()
for variant in (5,):
    got = torch.zeros_like(ref)
    run(variant, 0, x0, cnt, got)
    torch.cuda.synchronize()
    eq = bool(torch.equal(ref, got))
    ok_all &= eq
    print(json.dumps({"probe": "bit_exact_vs_throughput_kernel", "variant": variant, "count": cnt, "equal": eq}), flush=True)

e = torch.empty((cnt, 1025), dtype=torch.int32, device="cuda")
for variant, g, sizes in ((5, 0, (1, 8, 33, 37, 70, 74, 111, 148)), (40, 1, (1, 74, 148, 296))):
    for n in sizes:
        best, med = timed(lambda: run(variant, g, x0, n, e))
        print(json.dumps({"probe": "blind_rotate_latency", "variant": variant, "count": n, "ms_best": round(best, 4), "ms_med": round(med, 4)}), flush=True)

cf._lib.check(L.b2fhe_set_br_variant(0))
cf._lib.check(L.b2fhe_set_gates_per_cta(0))
cf._lib.check(L.b2fhe_set_latency_mode(1))
a = torch.from_numpy(orc.encrypt(keys, np.ones(148, np.uint8), 0).view(np.int16)).cuda()
o = torch.empty_like(a)
for n in (1, 8, 33, 70, 148):
    best, med = timed(lambda: cf._lib.check(L.b2fhe_gate_batch(0, None, 0, 0, o.data_ptr(), a.data_ptr(), a.data_ptr(), None, n)))
    print(json.dumps({"probe": "nand_gate_latency_mode", "count": n, "ms_best": round(best, 4), "ms_med": round(med, 4)}), flush=True)
print(json.dumps({"probe": "summary", "all_equal": ok_all}), flush=True)
cf.CleanUp()
