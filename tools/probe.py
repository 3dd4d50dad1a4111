"""Quick device-side timing probe (CUDA events) of the individual kernels; prints JSON lines.
Synthetic random keys and ciphertexts: timing does not depend on the values."""
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


for mode in (0, 1):
    g = C.c_double()
    cf._lib.check(L.b2fhe_int_peak(0, mode, C.byref(g)))
    print(json.dumps({"probe": "int_peak", "mode": mode, "giga_per_s": g.value}), flush=True)

cnt = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
x0 = torch.from_numpy(rng.integers(0, 2**16, (cnt, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
x1 = torch.from_numpy(rng.integers(0, 2**16, (cnt, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
e = torch.empty((cnt, 1025), dtype=torch.int32, device="cuda")
out = torch.empty_like(x0)
for variant, gs in ((32, (6,)), (16, (3, 4, 5, 6))):
    cf._lib.check(L.b2fhe_set_br_variant(variant))
    for g in gs:
        cf._lib.check(L.b2fhe_set_gates_per_cta(g))
        best, med = timed(lambda: cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, e.data_ptr(), x0.data_ptr(), 1 << 29, cnt)))
        print(json.dumps({"probe": "blind_rotate", "variant": variant, "gates_per_cta": g, "count": cnt, "ms_best": best,
                          "ms_med": med, "gates_per_s": cnt / best * 1e3}), flush=True)
cf._lib.check(L.b2fhe_set_br_variant(0))
cf._lib.check(L.b2fhe_set_gates_per_cta(0))
best, med = timed(lambda: cf._lib.check(L.b2fhe_keyswitch_batch(0, None, out.data_ptr(), e.data_ptr(), cnt)))
print(json.dumps({"probe": "keyswitch", "count": cnt, "ms_best": best, "gates_per_s": cnt / best * 1e3}), flush=True)
best, med = timed(lambda: cf._lib.check(L.b2fhe_gate_batch(0, None, 0, 0, out.data_ptr(), x0.data_ptr(), x1.data_ptr(), None, cnt)))
print(json.dumps({"probe": "nand_lvl0", "count": cnt, "ms_best": best, "gates_per_s": cnt / best * 1e3}), flush=True)
for small in (1, 148, 296, 444, 592, 148 * 6):
    for variant in (1, 16):
        cf._lib.check(L.b2fhe_set_br_variant(variant))
        best, med = timed(lambda: cf._lib.check(L.b2fhe_gate_batch(0, None, 0, 0, out.data_ptr(), x0.data_ptr(), x1.data_ptr(), None, small)))
        print(json.dumps({"probe": "nand_lvl0", "variant": variant, "count": small, "ms_best": best, "gates_per_s": small / best * 1e3}), flush=True)
cf._lib.check(L.b2fhe_set_br_variant(0))

npoly = 65536
a = torch.randint(-2**31, 2**31 - 1, (npoly, 1024), dtype=torch.int32, device="cuda")
b = torch.randint(-128, 128, (npoly, 1024), dtype=torch.int32, device="cuda")
o = torch.empty_like(a)
nt = torch.empty((npoly, 2, 1024), dtype=torch.int32, device="cuda")
for name, fn, modmul in (
        ("polymul", lambda: cf._lib.check(L.b2fhe_polymul1024_batch(0, None, o.data_ptr(), a.data_ptr(), b.data_ptr(), npoly)), 17408),
        ("ntt", lambda: cf._lib.check(L.b2fhe_ntt1024_batch(0, None, nt.data_ptr(), a.data_ptr(), npoly)), 5120),
        ("intt", lambda: cf._lib.check(L.b2fhe_intt1024_batch(0, None, o.data_ptr(), nt.data_ptr(), npoly)), 6144)):
    best, med = timed(fn, reps=5, warm=2)
    print(json.dumps({"probe": name, "count": npoly, "ms_best": best, "polys_per_s": npoly / best * 1e3}), flush=True)
cf.CleanUp()
