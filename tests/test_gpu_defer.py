"""GPU parity for the scattered-pointer batch (b2fhe_gate_batch_ptrs) and the deferred
per-gate queue (b2fhe_defer_*): how the reference's one-gate-per-call API
(src/cufhe_gates_gpu.cu:148-665, test/test_gate_gpu.cc:68-83) gets batched execution.
Bit-exact against the oracle, including the dependency cases a circuit issued gate by
gate produces (read-after-write, write-after-write, in-place, write-after-read)."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import orc

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int16 if a.dtype == np.uint16 else np.int32)).cuda()


def _np(t, level):
    return t.cpu().numpy().view(np.uint16 if level == 0 else np.uint32)


def _ptrs(rows):
    return (C.c_void_p * len(rows))(*[r.data_ptr() if r is not None else None for r in rows])


@pytest.mark.parametrize("level", [0, 1])
def test_gate_batch_ptrs(engine, keys, level):
    """Gates whose operands live in separately allocated buffers, mixed opcodes, an output
    aliasing an input: one gather -> batched path -> scatter."""
    cf = engine
    L = cf._lib.lib()
    cnt = 37
    ops = (np.arange(cnt) % 14).astype(np.uint8)
    rng = np.random.default_rng(11 + level)
    b = [rng.integers(0, 2, cnt) for _ in range(3)]
    x = [orc.encrypt(keys, b[k], level, base_index=61000 + 1000 * k) for k in range(3)]
    # every ciphertext in its own allocation, in shuffled order
    bufs = [[_dev(x[k][i].copy()) for i in range(cnt)] for k in range(3)]
    outs = [torch.zeros_like(bufs[0][0]) for _ in range(cnt)]
    for i in range(0, cnt, 5):
        outs[i] = bufs[0][i]                      # in place
    ref = orc.gate_batch(keys, 0, level, x[0], x[1], x[2], ops=ops)
    cf._lib.check(L.b2fhe_gate_batch_ptrs(0, None, ops.ctypes.data, level, _ptrs(outs), _ptrs(bufs[0]),
                                          _ptrs(bufs[1]), _ptrs(bufs[2]), cnt))
    cf.Synchronize()
    got = np.stack([_np(o, level) for o in outs])
    assert np.array_equal(got, ref)


def test_gate_batch_ptrs_errors(engine):
    cf = engine
    L = cf._lib.lib()
    ops = np.zeros(2, np.uint8)
    nul = (C.c_void_p * 2)(None, None)
    assert L.b2fhe_gate_batch_ptrs(0, None, ops.ctypes.data, 0, nul, nul, nul, None, 2) != 0     # null operands
    assert L.b2fhe_gate_batch_ptrs(0, None, ops.ctypes.data, 3, nul, nul, nul, None, 2) != 0     # bad level
    assert L.b2fhe_gate_batch_ptrs(0, None, None, 0, None, None, None, None, 0) == 0             # empty batch
    # a two-input gate with a null second operand / a Mux with a null third operand: rejected per gate
    d = torch.zeros((2, 637), dtype=torch.int16, device="cuda")
    ok = (C.c_void_p * 2)(d[0].data_ptr(), d[1].data_ptr())
    half = (C.c_void_p * 2)(d[0].data_ptr(), None)
    assert L.b2fhe_gate_batch_ptrs(0, None, ops.ctypes.data, 0, ok, ok, half, None, 2) == cf._lib.E_BAD_ARGUMENT
    assert L.b2fhe_gate_batch_ptrs(0, None, ops.ctypes.data, 0, ok, ok, None, None, 2) == cf._lib.E_BAD_ARGUMENT
    mux = np.full(2, orc.OP["MUX"], np.uint8)
    assert L.b2fhe_gate_batch_ptrs(0, None, mux.ctypes.data, 0, ok, ok, ok, half, 2) == cf._lib.E_BAD_ARGUMENT
    nots = np.full(2, orc.OP["NOT"], np.uint8)
    assert L.b2fhe_gate_batch_ptrs(0, None, nots.ctypes.data, 0, ok, ok, None, None, 2) == 0     # unary: in1 not needed
    torch.cuda.synchronize()
    assert L.b2fhe_defer_gate(0, None, 99, 0, C.c_void_p(8), C.c_void_p(8), None, None) != 0     # bad opcode
    assert L.b2fhe_defer_gate(0, None, orc.OP["MUX"], 0, C.c_void_p(8), C.c_void_p(8), C.c_void_p(8), None) != 0
    assert L.b2fhe_defer_pending(0) == 0


@pytest.mark.parametrize("level", [0, 1])
def test_deferred_queue_dependencies(engine, keys, level):
    """A 4-bit ripple-carry adder issued ONE GATE PER CALL, results written back in place:
    the queue has to levelise it; bytes equal the oracle's sequential evaluation."""
    cf = engine
    L = cf._lib.lib()
    OP = orc.OP
    nbits, lanes = 4, 5
    rng = np.random.default_rng(23 + level)
    A = rng.integers(0, 2, (lanes, nbits))
    B = rng.integers(0, 2, (lanes, nbits))
    words = 637 if level == 0 else 1025
    dt = np.uint16 if level == 0 else np.uint32
    host = {}       # name -> numpy ciphertext (oracle's view)
    dev = {}        # name -> device buffer

    def new(name, ct):
        host[name] = ct.copy()
        dev[name] = _dev(ct.copy())

    for l in range(lanes):
        ea = orc.encrypt(keys, A[l], level, base_index=70000 + 100 * l)
        eb = orc.encrypt(keys, B[l], level, base_index=75000 + 100 * l)
        for i in range(nbits):
            new(f"a{l}_{i}", ea[i]); new(f"b{l}_{i}", eb[i])
        for nm in ("c", "t", "u"):
            new(f"{nm}{l}", np.zeros(words, dt))
        for i in range(nbits):
            new(f"s{l}_{i}", np.zeros(words, dt))

    user = torch.cuda.Stream()
    torch.cuda.synchronize()

    def gate(op, out, i0, i1=None, i2=None):
        ins = [host[i0][None], host[i1][None] if i1 else None, host[i2][None] if i2 else None]
        host[out] = orc.gate_batch(keys, OP[op], level, *ins)[0]
        cf._lib.check(L.b2fhe_defer_gate(0, C.c_void_p(user.cuda_stream), OP[op], level, dev[out].data_ptr(),
                                         dev[i0].data_ptr(), dev[i1].data_ptr() if i1 else None,
                                         dev[i2].data_ptr() if i2 else None))

    for l in range(lanes):
        a = lambda i: f"a{l}_{i}"
        bb = lambda i: f"b{l}_{i}"
        s = lambda i: f"s{l}_{i}"
        c, t, u = f"c{l}", f"t{l}", f"u{l}"
        gate("XOR", s(0), a(0), bb(0))
        gate("AND", c, a(0), bb(0))
        for i in range(1, nbits):
            gate("XOR", t, a(i), bb(i))          # t rewritten every bit: write-after-read + write-after-write
            gate("XOR", s(i), t, c)
            gate("AND", u, t, c)
            gate("AND", t, a(i), bb(i))          # reuses t while u's producer read it
            gate("OR", c, u, t)                  # carry updated in place
        gate("NOT", a(0), a(0))                  # in-place on an input read by level-0 gates
        gate("MUX", bb(0), c, s(0), a(0))
    n_gates = lanes * (2 + 5 * (nbits - 1) + 2)
    assert L.b2fhe_defer_pending(0) == n_gates
    # result copies queued behind the gates
    out_host = {nm: torch.empty(words, dtype=dev[nm].dtype).pin_memory() for nm in dev}
    for nm in dev:
        cf._lib.check(L.b2fhe_defer_d2h(0, out_host[nm].data_ptr(), dev[nm].data_ptr(), words * (2 if level == 0 else 4)))
    busy = C.c_int(1)
    cf._lib.check(L.b2fhe_defer_query(0, C.byref(busy)))       # starts the queue
    cf.Synchronize()
    assert L.b2fhe_defer_pending(0) == 0
    cf._lib.check(L.b2fhe_defer_query(0, C.byref(busy)))
    assert busy.value == 0
    for nm in dev:
        assert np.array_equal(_np(out_host[nm], level), host[nm]), nm
    for l in range(lanes):
        va = sum(int(A[l][i]) << i for i in range(nbits)); vb = sum(int(B[l][i]) << i for i in range(nbits))
        bits = orc.decrypt(keys, np.stack([host[f"s{l}_{i}"] for i in range(nbits)] + [host[f"c{l}"]]), level)
        assert sum(int(v) << i for i, v in enumerate(bits)) == va + vb


def test_fence_orders_immediate_calls_after_queue(engine, keys):
    """An immediate batched call on a stream sees the results of gates queued before it."""
    cf = engine
    L = cf._lib.lib()
    x0 = orc.encrypt(keys, [1, 0, 1], 0, base_index=81000)
    x1 = orc.encrypt(keys, [1, 1, 0], 0, base_index=82000)
    d0, d1 = _dev(x0), _dev(x1)
    mid = torch.zeros_like(d0)
    st = torch.cuda.Stream()
    torch.cuda.synchronize()
    sp = C.c_void_p(st.cuda_stream)
    for i in range(3):
        cf._lib.check(L.b2fhe_defer_gate(0, sp, orc.OP["NAND"], 0, mid[i].data_ptr(), d0[i].data_ptr(), d1[i].data_ptr(), None))
    out = torch.zeros_like(d0)
    cf._lib.check(L.b2fhe_defer_fence(0, sp))
    cf._lib.check(L.b2fhe_gate_batch(0, sp, orc.OP["XOR"], 0, out.data_ptr(), mid.data_ptr(), d1.data_ptr(), None, 3))
    cf.Synchronize()
    m = orc.gate_batch(keys, orc.OP["NAND"], 0, x0, x1)
    assert np.array_equal(_np(mid, 0), m)
    assert np.array_equal(_np(out, 0), orc.gate_batch(keys, orc.OP["XOR"], 0, m, x1))


@pytest.mark.parametrize("level", [0, 1])
def test_deferred_copy_variant_chain(engine, keys, level):
    """What Nand(c, a, b, st); Nand(e, c, d, st) of the C++ shim queue: operand uploads and
    result copy-backs ride with the queue as staged transfers; the upload of `c` (whose host
    copy is still stale when the second call is made) must not clobber the first gate's result,
    neither in the same flush nor after the first gate was flushed on its own."""
    cf = engine
    L = cf._lib.lib()
    NAND = orc.OP["NAND"]
    words = 637 if level == 0 else 1025
    nbytes = words * (2 if level == 0 else 4)
    tdt = torch.int16 if level == 0 else torch.int32
    x = [orc.encrypt(keys, [1, 0, 1, 1], level, base_index=91000 + 1000 * k) for k in range(3)]     # a, b, d
    want_c = orc.gate_batch(keys, NAND, level, x[0], x[1])
    want_e = orc.gate_batch(keys, NAND, level, want_c, x[2])
    for flush_between in (False, True):
        host = {nm: [torch.from_numpy(x[k][i].copy().view(np.int16 if level == 0 else np.int32)).pin_memory() for i in range(4)]
                for k, nm in enumerate("abd")}
        host["c"] = [torch.full((words,), 0x5A5A if level == 0 else 0x5A5A5A5A, dtype=tdt).pin_memory() for _ in range(4)]
        host["e"] = [torch.zeros(words, dtype=tdt).pin_memory() for _ in range(4)]
        dev = {nm: [torch.zeros(words, dtype=tdt, device="cuda") for _ in range(4)] for nm in "abcde"}
        torch.cuda.synchronize()

        def call(out, i0, i1, i):
            for nm in (i0, i1):
                cf._lib.check(L.b2fhe_defer_h2d(0, dev[nm][i].data_ptr(), host[nm][i].data_ptr(), nbytes))
            cf._lib.check(L.b2fhe_defer_gate(0, None, NAND, level, dev[out][i].data_ptr(), dev[i0][i].data_ptr(),
                                             dev[i1][i].data_ptr(), None))
            cf._lib.check(L.b2fhe_defer_d2h(0, host[out][i].data_ptr(), dev[out][i].data_ptr(), nbytes))

        for i in range(4):
            call("c", "a", "b", i)
        if flush_between:
            cf._lib.check(L.b2fhe_defer_flush(0))
        for i in range(4):
            call("e", "c", "d", i)
        cf.Synchronize()
        got_c = np.stack([_np(t, level) for t in host["c"]])
        got_e = np.stack([_np(t, level) for t in host["e"]])
        assert np.array_equal(got_c, want_c), flush_between
        assert np.array_equal(got_e, want_e), flush_between
    # after Synchronize the host copies are authoritative again: a fresh upload is honoured
    c_dev = dev["c"][0]
    fresh = torch.from_numpy(x[0][1].copy().view(np.int16 if level == 0 else np.int32)).pin_memory()
    out_h = torch.zeros(words, dtype=tdt).pin_memory()
    cf._lib.check(L.b2fhe_defer_h2d(0, c_dev.data_ptr(), fresh.data_ptr(), nbytes))
    cf._lib.check(L.b2fhe_defer_gate(0, None, NAND, level, dev["e"][0].data_ptr(), c_dev.data_ptr(), dev["d"][0].data_ptr(), None))
    cf._lib.check(L.b2fhe_defer_d2h(0, out_h.data_ptr(), dev["e"][0].data_ptr(), nbytes))
    cf.Synchronize()
    assert np.array_equal(_np(out_h, level), orc.gate_batch(keys, NAND, level, x[0][1:2], x[2][0:1])[0])


def test_deferred_copy_back_is_a_snapshot(engine, keys):
    """Nand(c, a, b, st) followed by gNot(c, c, st): the host copy queued with the NAND holds the
    NAND (the reference copies it back before the NOT runs on that stream), a second copy-back
    queued after the NOT holds the NOT."""
    cf = engine
    L = cf._lib.lib()
    x0 = orc.encrypt(keys, [1, 1], 0, base_index=93000)
    x1 = orc.encrypt(keys, [1, 0], 0, base_index=94000)
    d0, d1 = _dev(x0), _dev(x1)
    c = torch.zeros_like(d0)
    h1 = [torch.zeros(637, dtype=torch.int16).pin_memory() for _ in range(2)]
    h2 = [torch.zeros(637, dtype=torch.int16).pin_memory() for _ in range(2)]
    torch.cuda.synchronize()
    for i in range(2):
        cf._lib.check(L.b2fhe_defer_gate(0, None, orc.OP["NAND"], 0, c[i].data_ptr(), d0[i].data_ptr(), d1[i].data_ptr(), None))
        cf._lib.check(L.b2fhe_defer_d2h(0, h1[i].data_ptr(), c[i].data_ptr(), 1274))
        cf._lib.check(L.b2fhe_defer_gate(0, None, orc.OP["NOT"], 0, c[i].data_ptr(), c[i].data_ptr(), None, None))
        cf._lib.check(L.b2fhe_defer_d2h(0, h2[i].data_ptr(), c[i].data_ptr(), 1274))
    cf.Synchronize()
    nand = orc.gate_batch(keys, orc.OP["NAND"], 0, x0, x1)
    assert np.array_equal(np.stack([_np(t, 0) for t in h1]), nand)
    assert np.array_equal(np.stack([_np(t, 0) for t in h2]), orc.gate_batch(keys, orc.OP["NOT"], 0, nand))
