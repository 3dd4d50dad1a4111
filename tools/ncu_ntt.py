"""Stand-alone NTT + INTT launches at batch 2^16 for ncu: python tools/ncu_ntt.py"""
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
n = 1 << 16
a = torch.from_numpy(rng.integers(0, 2**32, (n, 1024), dtype=np.uint64).astype(np.uint32).view(np.int32)).cuda()
nt = torch.empty((n, 2, 1024), dtype=torch.int32, device="cuda")
o = torch.empty((n, 1024), dtype=torch.int32, device="cuda")
for _ in range(2):
    cf._lib.check(L.b2fhe_ntt1024_batch(0, None, nt.data_ptr(), a.data_ptr(), n))
    cf._lib.check(L.b2fhe_intt1024_batch(0, None, o.data_ptr(), nt.data_ptr(), n))
torch.cuda.synchronize()
assert torch.equal(o, a)
cf.CleanUp()
