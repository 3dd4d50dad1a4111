"""One blind-rotation launch for ncu: python tools/ncu_br.py <variant> <gates_per_cta> <count>"""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import cufhe_b200 as cf  # noqa: E402
L = cf._lib.lib()
variant, g, cnt = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rng = np.random.default_rng(0)
bk = rng.integers(0, 2**32, cf.api.BK_RAW_WORDS, dtype=np.uint64).astype(np.uint32)
ksk = rng.integers(0, 2**16, cf.api.KSK_RAW_WORDS, dtype=np.uint64).astype(np.uint16)
cf.SetGPUNum(1)
cf.Initialize(bk, ksk)
cf._lib.check(L.b2fhe_set_br_variant(variant))
cf._lib.check(L.b2fhe_set_gates_per_cta(g))
x0 = torch.from_numpy(rng.integers(0, 2**16, (cnt, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
e = torch.empty((cnt, 1025), dtype=torch.int32, device="cuda")
for _ in range(2):
    cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, e.data_ptr(), x0.data_ptr(), 1 << 29, cnt))
torch.cuda.synchronize()
cf.CleanUp()
