"""Uniform lvl0 Mux over 16384 gates: ONE blind-rotation launch of 2 x count work items (55 full waves + 1
partial) instead of two launches of count (2 x (27 full + 1 partial)); timing with CUDA events."""
import json, sys
import numpy as np, torch
sys.path.insert(0, ".")
import cufhe_b200 as cf
L = cf._lib.lib()
rng = np.random.default_rng(0)
bk = rng.integers(0, 2**32, cf.api.BK_RAW_WORDS, dtype=np.uint64).astype(np.uint32)
ksk = rng.integers(0, 2**16, cf.api.KSK_RAW_WORDS, dtype=np.uint64).astype(np.uint16)
cf.SetGPUNum(1); cf.Initialize(bk, ksk)
n = 16384
x = [torch.from_numpy(rng.integers(0, 2**16, (n, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda() for _ in range(3)]
o = torch.empty_like(x[0])
def t(fn, reps=3):
    fn(); torch.cuda.synchronize(); best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
    return best
mux = t(lambda: cf._lib.check(L.b2fhe_gate_batch(0, None, cf.OP["MUX"], 0, o.data_ptr(), x[0].data_ptr(), x[1].data_ptr(), x[2].data_ptr(), n)))
nand = t(lambda: cf._lib.check(L.b2fhe_gate_batch(0, None, cf.OP["NAND"], 0, o.data_ptr(), x[0].data_ptr(), x[1].data_ptr(), None, n)))
print(json.dumps({"probe": "mux_16384_one_launch", "ms": mux, "nand_16384_ms": nand, "two_launch_estimate_ms": 2 * nand - 5.1,
                  "bootstraps_per_s": 2 * n / mux * 1e3}))
cf.CleanUp()
