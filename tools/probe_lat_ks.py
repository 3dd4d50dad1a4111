"""NAND latency (warm L2 and flushed L2) with the default k_keyswitch_lat (ks variant 0: predicated loads of a whole
coefficient's rows in flight, 32 warps per gate up to 4 gates per SM) and its round-1 form (3: a branch per digit, 16 warps).  JSON lines."""
import json, sys
import numpy as np, torch
sys.path.insert(0, ".")
import cufhe_b200 as cf
L = cf._lib.lib()
rng = np.random.default_rng(0)
bk = rng.integers(0, 2**32, cf.api.BK_RAW_WORDS, dtype=np.uint64).astype(np.uint32)
ksk = rng.integers(0, 2**16, cf.api.KSK_RAW_WORDS, dtype=np.uint64).astype(np.uint16)
cf.SetGPUNum(1); cf.Initialize(bk, ksk)
a = torch.from_numpy(rng.integers(0, 2**16, (4736, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
o = torch.empty_like(a)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(n, fl, reps=9):
    ts = []
    for _ in range(reps + 2):
        if fl: flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); cf._lib.check(L.b2fhe_gate_batch(0, None, 0, 0, o.data_ptr(), a.data_ptr(), a.data_ptr(), None, n)); e1.record()
        torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ts = ts[2:]
    return round(min(ts), 4), round(sorted(ts)[len(ts) // 2], 4)
for rep in range(2):
    for v in (0, 3):
        cf._lib.check(L.b2fhe_set_ks_variant(v))
        for n in (1, 33, 70, 148, 592):
            for fl in (False, True):
                b, m = t(n, fl)
                print(json.dumps({"probe": "nand_latency", "ks_variant": v, "count": n, "l2_flushed": fl, "ms_best": b, "ms_med": m}), flush=True)
cf._lib.check(L.b2fhe_set_ks_variant(0))
cf.CleanUp()
