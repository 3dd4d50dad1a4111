"""A/B of the k_blind_rotate16 key-multiply variants (16 = one row at a time, 17 = row pairs share
a Montgomery reduction, 18 = ... and the pair's digit transforms run together): outputs compared
bit for bit with variant 16, then CUDA-event timings.  Prints JSON lines."""
import ctypes as C
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import cufhe_b200 as cf  # noqa: E402

L = cf._lib.lib()
rng = np.random.default_rng(0)
bk = rng.integers(0, 2**32, cf.api.BK_RAW_WORDS, dtype=np.uint64).astype(np.uint32)
ksk = rng.integers(0, 2**16, cf.api.KSK_RAW_WORDS, dtype=np.uint64).astype(np.uint16)
cf.SetGPUNum(1)
cf.Initialize(bk, ksk)


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


cnt = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
x0 = torch.from_numpy(rng.integers(0, 2**16, (cnt, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()


def run(variant, g, n, dst):
    cf._lib.check(L.b2fhe_set_br_variant(variant))
    cf._lib.check(L.b2fhe_set_gates_per_cta(g))
    cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, dst.data_ptr(), x0.data_ptr(), 1 << 29, n))


nchk = min(cnt, 1200)
ref = torch.zeros((nchk, 1025), dtype=torch.int32, device="cuda")
run(16, 4, nchk, ref)
torch.cuda.synchronize()
for variant in (19,):
    for g in (1, 2, 3, 4):
        got = torch.zeros_like(ref)
        run(variant, g, nchk, got)
        torch.cuda.synchronize()
        print(json.dumps({"probe": "bit_exact_vs_16", "variant": variant, "gates_per_cta": g, "count": nchk,
                          "equal": bool(torch.equal(ref, got))}), flush=True)

e = torch.empty((cnt, 1025), dtype=torch.int32, device="cuda")
for variant, gs in ((16, (4,)), (19, (1, 2, 3, 4)),):
    for g in gs:
        best, med = timed(lambda: run(variant, g, cnt, e))
        print(json.dumps({"probe": "blind_rotate", "variant": variant, "gates_per_cta": g, "count": cnt, "ms_best": best,
                          "ms_med": med, "gates_per_s": cnt / best * 1e3}), flush=True)
# default path (variant 0, gates per CTA 0) with and without the separate launch of the last partial wave
for tail in (-1, -2, -1, -2):
    cf._lib.check(L.b2fhe_set_br_variant(tail))
    for n in (cnt, 592 * 5 + 100, 592 + 300):
        best, med = timed(lambda: run(0, 0, n, e))
        print(json.dumps({"probe": "blind_rotate_default", "tail_split": tail == -2, "count": n, "ms_best": best, "ms_med": med,
                          "gates_per_s": n / best * 1e3}), flush=True)
ref = torch.zeros((cnt, 1025), dtype=torch.int32, device="cuda")
got = torch.zeros_like(ref)
cf._lib.check(L.b2fhe_set_br_variant(-1))
run(0, 0, cnt, ref)
cf._lib.check(L.b2fhe_set_br_variant(-2))
run(0, 0, cnt, got)
torch.cuda.synchronize()
print(json.dumps({"probe": "tail_split_bit_exact", "count": cnt, "equal": bool(torch.equal(ref, got))}), flush=True)
cf._lib.check(L.b2fhe_set_br_variant(0))
cf._lib.check(L.b2fhe_set_gates_per_cta(0))
cf.CleanUp()
