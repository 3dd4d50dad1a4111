"""GPU test against the REAL reference CPU library: the CUDA engine is initialised with a
TFHEpp-generated EvalKey (raw arrays, exactly what cufhe::Initialize(ek) uploads), fed
TFHEpp-encrypted inputs, and its outputs are (a) byte-equal to the oracle on those keys and
(b) decrypted by TFHEpp to the truth table and to TFHEpp's own HomNAND result.
Runs last (file name) because it re-initialises the engine with different keys."""
import numpy as np
import pytest

from oracle import orc, tfhepp

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not tfhepp.available(), reason="oracle/_ref not built (needs /root/reference)")]


def test_cuda_engine_on_tfhepp_keys():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import cufhe_b200 as cf
    cf.CleanUp()
    tk = tfhepp.Keys()
    cf.SetGPUNum(1)
    cf.Initialize(tk.bk, tk.ksk)
    try:
        st = cf.Stream(0).Create()
        rng = np.random.default_rng(11)
        for level in (0, 1):
            cnt = 24
            b = [rng.integers(0, 2, cnt).astype(np.uint8) for _ in range(3)]
            x = [tfhepp.encrypt(b[k], level) for k in range(3)]
            ops = (np.arange(cnt) % 14).astype(np.uint8)
            cts = [cf.Ctxt(cnt, level).set(v) for v in x]
            out = cf.Ctxt(cnt, level)
            for c in cts:
                cf.CtxtCopyH2D(c, st)
            cf.gGateMixed(ops, out, cts[0], cts[1], cts[2], st)
            cf.CtxtCopyD2H(out, st)
            cf.Synchronize()
            assert np.array_equal(out.tlwehost, orc.gate_batch(tk, 0, level, x[0], x[1], x[2], ops=ops))
            dec = tfhepp.decrypt(out.tlwehost, level)
            assert list(dec) == [orc.truth(ops[i], b[0][i], b[1][i], b[2][i]) for i in range(cnt)]
        # NAND lvl0 vs TFHEpp's own gate
        cnt = 64
        b0, b1 = rng.integers(0, 2, cnt).astype(np.uint8), rng.integers(0, 2, cnt).astype(np.uint8)
        x0, x1 = tfhepp.encrypt(b0, 0), tfhepp.encrypt(b1, 0)
        c0, c1, out = cf.Ctxt(cnt, 0).set(x0), cf.Ctxt(cnt, 0).set(x1), cf.Ctxt(cnt, 0)
        cf.Nand(out, c0, c1, st)
        cf.Synchronize()
        ours = tfhepp.decrypt(out.tlwehost, 0)
        theirs = tfhepp.decrypt(tfhepp.homnand_lvl0(x0, x1), 0)
        assert list(ours) == list(theirs) == list(1 - b0 * b1)
    finally:
        cf.CleanUp()
