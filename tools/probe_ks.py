"""A/B of the key-switch kernels on a large batch: k_keyswitch (one warp per gate, free-running) vs
k_keyswitch_step<W, SYNCJ> (W gates per CTA in step).  Outputs compared bit for bit.  JSON lines."""
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
cnt = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
e = torch.randint(-2**31, 2**31 - 1, (cnt, 1025), dtype=torch.int32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def run(variant, dst):
    cf._lib.check(L.b2fhe_set_ks_variant(variant))
    cf._lib.check(L.b2fhe_keyswitch_batch(0, None, dst.data_ptr(), e.data_ptr(), cnt))


def timed(variant, dst, reps=5):
    ts = []
    for _ in range(reps + 1):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(variant, dst)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts[1:]), sorted(ts[1:])[len(ts[1:]) // 2]


ref = torch.zeros((cnt, 637), dtype=torch.int16, device="cuda")
run(2, ref)
torch.cuda.synchronize()
for v in [int(a) for a in sys.argv[2:]] or (2, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 0, 2):
    got = torch.zeros_like(ref)
    best, med = timed(v, got)
    print(json.dumps({"probe": "keyswitch", "ks_variant": v, "count": cnt, "ms_best": round(best, 4), "ms_med": round(med, 4),
                      "equal": bool(torch.equal(ref, got))}), flush=True)
cf._lib.check(L.b2fhe_set_ks_variant(0))
cf.CleanUp()
