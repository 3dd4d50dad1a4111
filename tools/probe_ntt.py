"""Timing + round-trip check of the stand-alone transform entry points (BASELINE config 2, batch 2^16)."""
import json, sys
import numpy as np, torch
sys.path.insert(0, ".")
import cufhe_b200 as cf
L = cf._lib.lib()
rng = np.random.default_rng(0)
bk = rng.integers(0, 2**32, cf.api.BK_RAW_WORDS, dtype=np.uint64).astype(np.uint32)
ksk = rng.integers(0, 2**16, cf.api.KSK_RAW_WORDS, dtype=np.uint64).astype(np.uint16)
cf.SetGPUNum(1); cf.Initialize(bk, ksk)
n = 1 << 16
a = torch.from_numpy(rng.integers(0, 2**32, (n, 1024), dtype=np.uint64).astype(np.uint32).view(np.int32)).cuda()
b = torch.from_numpy(rng.integers(0, 1 << 18, (n, 1024), dtype=np.int64).astype(np.int32)).cuda()
nt = torch.empty((n, 2, 1024), dtype=torch.int32, device="cuda"); o = torch.empty((n, 1024), dtype=torch.int32, device="cuda")
def t(fn, reps=5):
    fn(); torch.cuda.synchronize(); best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
    return best
for pf in (-3, -4):
  cf._lib.check(L.b2fhe_set_br_variant(pf))
  for name, fn in (("ntt", lambda: cf._lib.check(L.b2fhe_ntt1024_batch(0, None, nt.data_ptr(), a.data_ptr(), n))),
                   ("intt", lambda: cf._lib.check(L.b2fhe_intt1024_batch(0, None, o.data_ptr(), nt.data_ptr(), n))),
                   ("polymul2", lambda: cf._lib.check(L.b2fhe_polymul1024_batch(0, None, o.data_ptr(), a.data_ptr(), b.data_ptr(), n))),
                   ("polymul3_exact", lambda: cf._lib.check(L.b2fhe_polymul1024_exact_batch(0, None, o.data_ptr(), a.data_ptr(), b.data_ptr(), n)))):
    ms = t(fn)
    print(json.dumps({"probe": name, "input_prefetch": pf == -4, "batch": n, "ms": ms, "gb_per_s": 12288 * n / ms * 1e-6}), flush=True)
cf._lib.check(L.b2fhe_ntt1024_batch(0, None, nt.data_ptr(), a.data_ptr(), n))
cf._lib.check(L.b2fhe_intt1024_batch(0, None, o.data_ptr(), nt.data_ptr(), n))
torch.cuda.synchronize()
print(json.dumps({"probe": "roundtrip_equal", "equal": bool(torch.equal(o, a))}))
cf.CleanUp()
