"""Latency probe for small batches: blind rotation (each kernel) and key switch separately."""
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


def timed(fn, reps=5, warm=2):
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
    return min(ts)


cnt_max = 2048
x0 = torch.from_numpy(rng.integers(0, 2**16, (cnt_max, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
e = torch.empty((cnt_max, 1025), dtype=torch.int32, device="cuda")
e.random_()
out = torch.empty_like(x0)
for cnt in [int(a) for a in sys.argv[1:]] or [1, 8, 16, 24, 32, 37, 64, 148]:
    row = {"count": cnt}
    for variant in (5, 4, 2, 1, 16, 19):
        cf._lib.check(L.b2fhe_set_br_variant(variant))
        row[f"br_v{variant}_ms"] = round(timed(lambda: cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, e.data_ptr(), x0.data_ptr(), 1 << 29, cnt))), 4)
    cf._lib.check(L.b2fhe_set_br_variant(19))
    for g in (1, 2):
        cf._lib.check(L.b2fhe_set_gates_per_cta(g))
        row[f"br_v19_g{g}_ms"] = round(timed(lambda: cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, e.data_ptr(), x0.data_ptr(), 1 << 29, cnt))), 4)
    cf._lib.check(L.b2fhe_set_gates_per_cta(0))
    cf._lib.check(L.b2fhe_set_br_variant(0))
    row["br_auto_ms"] = round(timed(lambda: cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, e.data_ptr(), x0.data_ptr(), 1 << 29, cnt))), 4)
    for kv in (1, 2):
        cf._lib.check(L.b2fhe_set_ks_variant(kv))
        row[f"ks_v{kv}_ms"] = round(timed(lambda: cf._lib.check(L.b2fhe_keyswitch_batch(0, None, out.data_ptr(), e.data_ptr(), cnt))), 4)
    cf._lib.check(L.b2fhe_set_ks_variant(0))
    row["nand_ms"] = round(timed(lambda: cf._lib.check(L.b2fhe_gate_batch(0, None, 0, 0, out.data_ptr(), x0.data_ptr(), x0.data_ptr(), None, cnt))), 4)
    print(json.dumps(row), flush=True)
cf.CleanUp()
