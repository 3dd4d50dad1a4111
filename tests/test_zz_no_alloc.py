"""b2fhe_gate_batch_noalloc (no counterpart in the reference):
the gate batch whose stage scratch is a fixed device array instead of a pool
allocation.  Called through the C ABI, two calls in a row, bytes compared with the existing entry
point b2fhe_gate_batch (tolerance 0: the path is integer-only) and with the oracle.

Collected last on purpose (file name).
"""
import numpy as np
import pytest
import torch

from oracle import orc

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int16)).cuda()


def test_no_alloc_gate_batch_matches_gate_batch_twice(engine, keys):
    cf = engine
    L = cf._lib.lib()
    if L.b2fhe_num_devices() < 1:            # an earlier test file may have cleaned up
        cf.SetGPUNum(1)
        cf.Initialize(keys.bk, keys.ksk)
    assert L.b2fhe_noalloc_max_gates() == 2048
    rng = np.random.default_rng(7)
    # second call: another op, fewer gates than the first (stale slots of call 1 stay in the array)
    for call, (opname, cnt) in enumerate((("NAND", 700), ("XOR", 37))):
        op = orc.OP[opname]
        b0, b1 = rng.integers(0, 2, cnt), rng.integers(0, 2, cnt)
        x0 = orc.encrypt(keys, b0, 0, base_index=5000 + 1000 * call)
        x1 = orc.encrypt(keys, b1, 0, base_index=9000 + 1000 * call)
        if call == 0:
            x0[3, :] = 0                     # all-zero ciphertext
            x1[4, :636] = 0xFFFF             # maximal mask values
        d0, d1 = _dev(x0), _dev(x1)
        want = torch.zeros_like(d0)
        got = torch.zeros_like(d0)
        cf._lib.check(L.b2fhe_gate_batch(0, None, op, 0, want.data_ptr(), d0.data_ptr(), d1.data_ptr(), None, cnt))
        cf._lib.check(L.b2fhe_gate_batch_noalloc(0, None, op, got.data_ptr(), d0.data_ptr(), d1.data_ptr(), cnt))
        torch.cuda.synchronize()
        w = want.cpu().numpy().view(np.uint16)
        g = got.cpu().numpy().view(np.uint16)
        assert np.array_equal(g, w), f"call {call} ({opname}): differs from b2fhe_gate_batch"
        nchk = 12                            # the oracle on a few gates: the checker is slow
        ref = orc.gate_batch(keys, op, 0, x0[:nchk], x1[:nchk])
        assert np.array_equal(g[:nchk], ref), f"call {call} ({opname}): differs from the oracle"
        dec = orc.decrypt(keys, g, 0)
        assert list(dec[5:]) == [orc.truth(op, int(b0[i]), int(b1[i]), 0) for i in range(5, cnt)]


def test_no_alloc_rejects_oversize_and_bad_ops(engine, keys):
    cf = engine
    L = cf._lib.lib()
    if L.b2fhe_num_devices() < 1:
        cf.SetGPUNum(1)
        cf.Initialize(keys.bk, keys.ksk)
    d = torch.zeros((4, 637), dtype=torch.int16, device="cuda")
    too_many = L.b2fhe_noalloc_max_gates() + 1
    assert L.b2fhe_gate_batch_noalloc(0, None, orc.OP["NAND"], d.data_ptr(), d.data_ptr(), d.data_ptr(), too_many) == cf._lib.E_TOO_LARGE
    assert b"exceeds the fixed scratch" in L.b2fhe_last_error_string()
    assert L.b2fhe_gate_batch_noalloc(0, None, orc.OP["MUX"], d.data_ptr(), d.data_ptr(), d.data_ptr(), 4) == cf._lib.E_BAD_ARGUMENT
    assert L.b2fhe_gate_batch_noalloc(0, None, orc.OP["NAND"], d.data_ptr(), d.data_ptr(), None, 4) == cf._lib.E_BAD_ARGUMENT
    assert L.b2fhe_gate_batch_noalloc(0, None, orc.OP["NAND"], d.data_ptr(), d.data_ptr(), d.data_ptr(), 0) == 0
