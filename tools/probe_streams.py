"""The reference's usage pattern (test/test_gate_gpu.cc, test_util.h:29-72): ONE gate per call,
many streams, H2D + gate + D2H per call.  Prints gates/s for per-gate NAND calls."""
import json
import sys
import time
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
nst = int(sys.argv[1]) if len(sys.argv) > 1 else 592
ngates = int(sys.argv[2]) if len(sys.argv) > 2 else 1184
mode = sys.argv[3] if len(sys.argv) > 3 else "copy"      # copy | resident | stage(br only) | defer (queued per-gate calls)
streams = [cf.Stream(0).Create() for _ in range(nst)]
rnd = rng.integers(0, 2**16, (ngates, 637), dtype=np.uint64).astype(np.uint16)
a = [cf.Ctxt(1, 0).set(rnd[i]) for i in range(ngates)]
b = [cf.Ctxt(1, 0).set(rnd[ngates - 1 - i]) for i in range(ngates)]
o = [cf.Ctxt(1, 0) for _ in range(ngates)]
e = [torch.empty((1, 1025), dtype=torch.int32, device="cuda") for _ in range(ngates if mode == "stage" else 0)]
for rep in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(ngates):
        if mode == "copy":
            cf.Nand(o[i], a[i], b[i], streams[i % nst])
        elif mode == "defer":
            st = streams[i % nst]
            cf._lib.check(L.b2fhe_defer_h2d(0, a[i].ptr(), a[i].tlwehost.ctypes.data, 637 * 2))
            cf._lib.check(L.b2fhe_defer_h2d(0, b[i].ptr(), b[i].tlwehost.ctypes.data, 637 * 2))
            cf._lib.check(L.b2fhe_defer_gate(0, st.st(), 0, 0, o[i].ptr(), a[i].ptr(), b[i].ptr(), None))
            cf._lib.check(L.b2fhe_defer_d2h(0, o[i].tlwehost.ctypes.data, o[i].ptr(), 637 * 2))
        elif mode == "resident":
            cf.gNand(o[i], a[i], b[i], streams[i % nst])
        else:
            cf._lib.check(L.b2fhe_blind_rotate_batch(0, streams[i % nst].st(), e[i].data_ptr(), a[i].ptr(), 1 << 29, 1))
    t1 = time.perf_counter()
    cf.Synchronize()
    t2 = time.perf_counter()
    print(json.dumps({"probe": "per_gate_calls", "mode": mode, "streams": nst, "gates": ngates, "issue_s": t1 - t0, "total_s": t2 - t0,
                      "gates_per_s": ngates / (t2 - t0), "ms_per_gate_throughput": (t2 - t0) * 1e3 / ngates}), flush=True)
cf.CleanUp()
