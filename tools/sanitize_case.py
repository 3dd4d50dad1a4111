"""Small launches of the synchronisation-heavy kernels for compute-sanitizer (one --tool per gpurun call):
k_blind_rotate_s<4> / <1> (named barriers + __syncwarp exchanges), k_blind_rotate_cl4 (st.async + mbarrier
across a 4-CTA cluster), k_keyswitch_step<16, 8> with a partial last CTA."""
import sys
import numpy as np, torch
sys.path.insert(0, ".")
import cufhe_b200 as cf
L = cf._lib.lib()
rng = np.random.default_rng(0)
bk = rng.integers(0, 2**32, cf.api.BK_RAW_WORDS, dtype=np.uint64).astype(np.uint32)
ksk = rng.integers(0, 2**16, cf.api.KSK_RAW_WORDS, dtype=np.uint64).astype(np.uint16)
cf.SetGPUNum(1); cf.Initialize(bk, ksk)
x = torch.from_numpy(rng.integers(0, 2**16, (8, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
e = torch.empty((21, 1025), dtype=torch.int32, device="cuda")
for variant, g, cnt in ((40, 4, 6), (40, 1, 2), (5, 0, 2)):
    cf._lib.check(L.b2fhe_set_br_variant(variant)); cf._lib.check(L.b2fhe_set_gates_per_cta(g))
    cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, e.data_ptr(), x.data_ptr(), 1 << 29, cnt))
    torch.cuda.synchronize(); print("ran blind rotation variant", variant, "gates/cta", g, "count", cnt, flush=True)
cf._lib.check(L.b2fhe_set_br_variant(0)); cf._lib.check(L.b2fhe_set_gates_per_cta(0))
y = torch.from_numpy(rng.integers(0, 2**32, (21, 1025), dtype=np.uint64).astype(np.uint32).view(np.int32)).cuda()
o = torch.empty((21, 637), dtype=torch.int16, device="cuda")
cf._lib.check(L.b2fhe_set_ks_variant(21))
cf._lib.check(L.b2fhe_keyswitch_batch(0, None, o.data_ptr(), y.data_ptr(), 21))
torch.cuda.synchronize(); print("ran k_keyswitch_step<16, 8> on 21 gates", flush=True)
cf._lib.check(L.b2fhe_set_ks_variant(0))
cf.CleanUp()
