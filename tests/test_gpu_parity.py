"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the
same seeded inputs -- bit-exact, as the hot path is integer-only.

Mirrors the reference's test strategy (test/test_gate_gpu.cc, test_util.h,
test_polynomial_mult_1024.cu, test_intensive.cc) but compares ciphertext BYTES, not
just decryptions.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import orc

pytestmark = pytest.mark.gpu

N = 1024


def dev_u32(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int32)).cuda()


def to_np(t, dtype):
    return t.cpu().numpy().view(dtype)


def run_gate(cf, op, level, ins, count):
    st = cf.Stream(0).Create()
    cts = [cf.Ctxt(count, level).set(x) if x is not None else None for x in ins]
    out = cf.Ctxt(count, level)
    for c in cts:
        if c is not None:
            cf.CtxtCopyH2D(c, st)
    cf.api._gate_device(op, out, cts[0], cts[1], cts[2], st)
    cf.CtxtCopyD2H(out, st)
    cf.Synchronize()
    return out.tlwehost.copy()


# ---------------------------------------------------------------- NTT layer ---
def test_ntt_roundtrip_and_polymul(engine):
    cf = engine
    L = cf._lib.lib()
    rng = np.random.default_rng(1)
    cnt = 300
    a = rng.integers(0, 2**32, size=(cnt, N), dtype=np.uint64).astype(np.uint32)
    da = dev_u32(a)
    dn = torch.empty((cnt, 2, N), dtype=torch.int32, device="cuda")
    db = torch.empty((cnt, N), dtype=torch.int32, device="cuda")
    cf._lib.check(L.b2fhe_ntt1024_batch(0, None, dn.data_ptr(), da.data_ptr(), cnt))
    cf._lib.check(L.b2fhe_intt1024_batch(0, None, db.data_ptr(), dn.data_ptr(), cnt))
    torch.cuda.synchronize()
    # INTT(NTT(x)) is the centred representative of x mod P, P ~ 2^54 > 2^32: exact
    assert np.array_equal(to_np(db, np.uint32), a)
    res = to_np(dn, np.uint32)
    assert res[:, 0].max() < 134215681 and res[:, 1].max() < 132120577

    # negacyclic product against the reference's schoolbook truth
    # (test/test_polynomial_mult_1024.cu:51-73), b = signed gadget digits and 9-bit values
    for lo, hi in ((-128, 128), (-512, 512)):
        b = rng.integers(lo, hi, size=(cnt, N)).astype(np.int32).view(np.uint32)
        dbb = dev_u32(b)
        dout = torch.empty((cnt, N), dtype=torch.int32, device="cuda")
        cf._lib.check(L.b2fhe_polymul1024_batch(0, None, dout.data_ptr(), da.data_ptr(), dbb.data_ptr(), cnt))
        torch.cuda.synchronize()
        got = to_np(dout, np.uint32)
        for i in range(0, cnt, 7):
            assert np.array_equal(got[i], orc.polymul(a[i], b[i], schoolbook=True)), f"poly {i}"


def test_ntt_input_prefetch_equals_plain_loads(engine):
    """The bulk-copy (TMA) input prefetch of b2fhe_ntt1024_batch / b2fhe_intt1024_batch against the plain-load
    kernels (b2fhe_set_br_variant(-3)): same bytes on a ragged batch long enough that every persistent team
    takes several polynomials (so the prefetched buffers are re-armed and re-used), on a batch of one, and
    from a source that is only 4-byte aligned (takes the plain loads)."""
    cf = engine
    L = cf._lib.lib()
    rng = np.random.default_rng(11)
    cnt = 5003
    a = rng.integers(0, 2**32, size=(cnt + 1, N), dtype=np.uint64).astype(np.uint32)
    da = dev_u32(a.reshape(-1))
    dsmall = dev_u32(rng.integers(-128, 128, size=(cnt, N)).astype(np.int32).view(np.uint32))
    outs = {}
    try:
        for pf in (-3, -4):
            cf._lib.check(L.b2fhe_set_br_variant(pf))
            dn = torch.zeros((cnt, 2, N), dtype=torch.int32, device="cuda")
            db = torch.zeros((cnt, N), dtype=torch.int32, device="cuda")
            d1 = torch.zeros((1, 2, N), dtype=torch.int32, device="cuda")
            dm = torch.zeros((7, 2, N), dtype=torch.int32, device="cuda")
            cf._lib.check(L.b2fhe_ntt1024_batch(0, None, dn.data_ptr(), da.data_ptr(), cnt))
            cf._lib.check(L.b2fhe_intt1024_batch(0, None, db.data_ptr(), dn.data_ptr(), cnt))
            cf._lib.check(L.b2fhe_ntt1024_batch(0, None, d1.data_ptr(), da.data_ptr() + 4096 * 17, 1))
            cf._lib.check(L.b2fhe_ntt1024_batch(0, None, dm.data_ptr(), da.data_ptr() + 4, 7))      # 4-byte aligned source
            # the same switch selects the CTA shape of the products (five / three products per CTA, or two)
            p2 = torch.zeros((cnt, N), dtype=torch.int32, device="cuda")
            p3 = torch.zeros((cnt, N), dtype=torch.int32, device="cuda")
            cf._lib.check(L.b2fhe_polymul1024_batch(0, None, p2.data_ptr(), da.data_ptr(), dsmall.data_ptr(), cnt))
            cf._lib.check(L.b2fhe_polymul1024_exact_batch(0, None, p3.data_ptr(), da.data_ptr(), da.data_ptr() + 4096 * 3, cnt - 3))
            pm = torch.zeros((9, N), dtype=torch.int32, device="cuda")
            cf._lib.check(L.b2fhe_polymul1024_batch(0, None, pm.data_ptr(), da.data_ptr() + 4, dsmall.data_ptr(), 9))      # 4-byte aligned operand
            torch.cuda.synchronize()
            outs[pf] = (to_np(dn, np.uint32), to_np(db, np.uint32), to_np(d1, np.uint32), to_np(dm, np.uint32), to_np(p2, np.uint32), to_np(p3, np.uint32),
                        to_np(pm, np.uint32))
    finally:
        cf._lib.check(L.b2fhe_set_br_variant(-4))
    for x, y in zip(outs[-3], outs[-4]):
        assert np.array_equal(x, y)
    assert np.array_equal(outs[-4][1], a[:cnt])                       # round trip
    assert np.array_equal(outs[-4][2][0], outs[-4][0][17])
    assert np.array_equal(outs[-4][4][5], orc.polymul(a[5], to_np(dsmall, np.uint32)[5], schoolbook=True))
    assert np.array_equal(outs[-4][5][cnt - 4], orc.polymul(a[cnt - 4], a[cnt - 1], schoolbook=True))


def test_polymul_edge_cases(engine):
    cf = engine
    L = cf._lib.lib()
    a = np.zeros((4, N), np.uint32)
    b = np.zeros((4, N), np.uint32)
    a[0, :] = 0xFFFFFFFF; b[0, :] = np.int32(-128).astype(np.uint32)     # extreme magnitudes
    a[1, 0] = 1; b[1, N - 1] = 1                                          # X^1023
    a[2, N - 1] = 0x80000000; b[2, N - 1] = 127                           # wrap with sign flip
    a[3, :] = np.arange(N, dtype=np.uint32) * 2654435761 & 0xFFFFFFFF; b[3, 5] = np.uint32(0xFFFFFFFF)  # * -X^5
    da, db = dev_u32(a), dev_u32(b)
    dout = torch.empty((4, N), dtype=torch.int32, device="cuda")
    cf._lib.check(L.b2fhe_polymul1024_batch(0, None, dout.data_ptr(), da.data_ptr(), db.data_ptr(), 4))
    cf._lib.check(L.b2fhe_polymul1024_batch(0, None, dout.data_ptr(), da.data_ptr(), db.data_ptr(), 0))  # empty batch
    torch.cuda.synchronize()
    got = to_np(dout, np.uint32)
    for i in range(4):
        assert np.array_equal(got[i], orc.polymul(a[i], b[i], schoolbook=True)), i


def test_polymul_exact_reference_domain_batch_2e16(engine):
    """BASELINE config 2 on the reference test's own input domain (test/test_polynomial_mult_1024.cu:38-39,
    161-163): a = any u32, b < 2^18, batch 2^16 -- the three-prime product must equal the schoolbook
    truth (test_polynomial_mult_1024.cu:51-73) with tolerance 0 (the reference allows 2) on a 256-polynomial
    subsample, plus full-range / extreme operands."""
    cf = engine
    L = cf._lib.lib()
    rng = np.random.default_rng(2)
    cnt = 1 << 16
    a = rng.integers(0, 2**32, size=(cnt, N), dtype=np.uint64).astype(np.uint32)
    b = rng.integers(0, 1 << 18, size=(cnt, N), dtype=np.uint64).astype(np.uint32)
    b[1] = rng.integers(0, 2**32, size=N, dtype=np.uint64).astype(np.uint32)        # any int32
    a[2], b[2] = 0xFFFFFFFF, 0x80000000                                              # extreme magnitudes
    a[3], b[3] = 0, 0x3FFFF
    da, db = dev_u32(a), dev_u32(b)
    dout = torch.empty((cnt, N), dtype=torch.int32, device="cuda")
    cf._lib.check(L.b2fhe_polymul1024_exact_batch(0, None, dout.data_ptr(), da.data_ptr(), db.data_ptr(), cnt))
    torch.cuda.synchronize()
    got = to_np(dout, np.uint32)
    idx = list(range(8)) + list(range(8, cnt, cnt // 248))[:248]
    for i in idx:
        assert np.array_equal(got[i], orc.polymul(a[i], b[i], schoolbook=True)), f"poly {i}"
    # the two-prime product is exact inside its documented bound |sum| < 2^52 (b2fhe.h) ...
    b9 = rng.integers(-512, 512, size=(64, N)).astype(np.int32).view(np.uint32)
    d9 = dev_u32(b9)
    o2 = torch.empty((64, N), dtype=torch.int32, device="cuda")
    o3 = torch.empty_like(o2)
    cf._lib.check(L.b2fhe_polymul1024_batch(0, None, o2.data_ptr(), da.data_ptr(), d9.data_ptr(), 64))
    cf._lib.check(L.b2fhe_polymul1024_exact_batch(0, None, o3.data_ptr(), da.data_ptr(), d9.data_ptr(), 64))
    torch.cuda.synchronize()
    assert torch.equal(o2, o3)
    # ... and NOT beyond it (which is why the exact entry point exists): 18-bit b overflows P ~ 2^53
    cf._lib.check(L.b2fhe_polymul1024_batch(0, None, o2.data_ptr(), da.data_ptr(), db.data_ptr(), 64))
    torch.cuda.synchronize()
    assert not np.array_equal(to_np(o2, np.uint32)[0], got[0])


# ------------------------------------------------------------------ stages ---
def test_blind_rotate_and_keyswitch_stages(engine, keys):
    cf = engine
    L = cf._lib.lib()
    cnt = 13          # ragged: not a multiple of any gates-per-CTA choice
    bits = np.arange(cnt) & 1
    c0 = orc.encrypt(keys, bits, 0)
    c0[3, :] = 0                      # all-zero ciphertext: every abar = 0, bbar = 2N
    c0[4, :636] = 0xFFFF              # maximal mask values
    d0 = torch.from_numpy(c0.view(np.int16)).cuda()
    d1 = torch.empty((cnt, 1025), dtype=torch.int32, device="cuda")
    refs = [orc.blind_rotate_extract(keys, c0[i]) for i in range(cnt)]
    # all blind-rotation kernels (5 = 4-CTA cluster kernel with st.async + mbarrier, 40 = the
    # throughput kernel k_blind_rotate_s, 0 = by batch size),
    # every gates-per-CTA instantiation
    for variant, gs in ((5, (0,)), (40, (0, 1, 2, 3, 4)), (0, (0,))):
        cf._lib.check(L.b2fhe_set_br_variant(variant))
        for g in gs:
            cf._lib.check(L.b2fhe_set_gates_per_cta(g))
            d1.zero_()
            cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, d1.data_ptr(), d0.data_ptr(), 1 << 29, cnt))
            torch.cuda.synchronize()
            got = to_np(d1, np.uint32)
            for i in range(cnt):
                assert np.array_equal(got[i], refs[i]), f"variant {variant}, gates/cta {g}, item {i}"
    cf._lib.check(L.b2fhe_set_gates_per_cta(0))
    cf._lib.check(L.b2fhe_set_br_variant(0))
    # key switch of those lvl1 outputs (+ an extreme input)
    x = got.copy()
    x[5, :] = 0xFFFFFFFF
    x[6, :] = 0
    dx = dev_u32(x)
    dks = torch.empty((cnt, 637), dtype=torch.int16, device="cuda")
    want = [orc.keyswitch(keys, x[i]) for i in range(cnt)]
    for ks_variant in (1, 2, 3, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 0):   # 16 warps per gate / one warp per gate / in-step CTAs (W x SYNCJ) / by batch size
        cf._lib.check(L.b2fhe_set_ks_variant(ks_variant))
        dks.zero_()
        cf._lib.check(L.b2fhe_keyswitch_batch(0, None, dks.data_ptr(), dx.data_ptr(), cnt))
        torch.cuda.synchronize()
        gk = to_np(dks, np.uint16)
        for i in range(cnt):
            assert np.array_equal(gk[i], want[i]), (ks_variant, i)


def test_schedule_jitter_is_bit_exact(engine, keys):
    """Schedule fuzzing in place of compute-sanitizer's racecheck / synccheck (closed on this pool): variants 41 / 6
    are k_blind_rotate_s / k_blind_rotate_cl4s compiled with a random 0..1 us sleep of every warp before and
    after every named barrier, __syncwarp, __syncthreads, mbarrier wait and DSMEM hand-over (clock-seeded, so
    every launch draws a different schedule).  A missing or misplaced synchronisation would let a warp read
    an exchange buffer / accumulator / receive buffer too early or overwrite it too late; the outputs must
    stay byte-identical to the oracle on every repetition."""
    cf = engine
    L = cf._lib.lib()
    cnt = 9
    c0 = orc.encrypt(keys, np.arange(cnt) & 1, 0, base_index=977)
    c0[2, :] = 0
    c0[3, :636] = 0xFFFF
    d0 = torch.from_numpy(c0.view(np.int16)).cuda()
    d1 = torch.empty((cnt, 1025), dtype=torch.int32, device="cuda")
    refs = np.stack([orc.blind_rotate_extract(keys, c0[i]) for i in range(cnt)])
    try:
        for variant, g in ((41, 4), (41, 1), (6, 0)):
            cf._lib.check(L.b2fhe_set_br_variant(variant))
            cf._lib.check(L.b2fhe_set_gates_per_cta(g))
            for rep in range(4):
                d1.zero_()
                cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, d1.data_ptr(), d0.data_ptr(), 1 << 29, cnt))
                torch.cuda.synchronize()
                assert np.array_equal(to_np(d1, np.uint32), refs), f"variant {variant}, gates/cta {g}, repetition {rep}"
    finally:
        cf._lib.check(L.b2fhe_set_gates_per_cta(0))
        cf._lib.check(L.b2fhe_set_br_variant(0))


def test_trlwe_level_entry_points(engine, keys):
    """BootstrapTLWE2TRLWE, SampleExtractAndKeySwitch, Refresh through the
    mirrored API, bytes against the oracle, ragged batch."""
    cf = engine
    cnt = 5
    bits = np.array([1, 0, 1, 1, 0], np.uint8)
    c0 = orc.encrypt(keys, bits, 0, base_index=4242)
    st = cf.Stream(0).Create()
    cin = cf.Ctxt(cnt, 0).set(c0)
    t1, t2 = cf.cuFHETRLWElvl1(cnt), cf.cuFHETRLWElvl1(cnt)
    out = cf.Ctxt(cnt, 0)
    cf.GateBootstrappingTLWE2TRLWElvl01NTT(t1, cin, st)
    cf.gRefresh(t2, t1, st)
    cf.gSampleExtractAndKeySwitch(out, t2, st)
    cf.CtxtCopyD2H(out, st)
    with torch.cuda.stream(st.torch_stream()):
        t2._host_t.copy_(t2.trlwedevice, non_blocking=True)
    cf.Synchronize()
    for i in range(cnt):
        w1 = orc.blind_rotate_trlwe(keys, c0[i])
        assert np.array_equal(t1.trlwehost[i], w1), i
        w2 = orc.refresh(keys, w1)
        assert np.array_equal(t2.trlwehost[i], w2), i
        assert np.array_equal(out.tlwehost[i], orc.sample_extract_keyswitch(keys, w2)), i
    assert list(orc.decrypt(keys, out.tlwehost, 0)) == list(bits)
    # Bootstrap with a run-time test-vector constant (src/bootstrap_gpu.cu:782-789), in place
    for mu in (1 << 29, 1 << 28, 0x12345678):
        cf.CtxtCopyH2D(cin, st)
        cf.gBootstrap(cin, cin, mu, st)
        cf.CtxtCopyD2H(cin, st)
        cf.Synchronize()
        for i in range(cnt):
            want = orc.keyswitch(keys, orc.blind_rotate_extract(keys, c0[i], mu=mu))
            assert np.array_equal(cin.tlwehost[i], want), (hex(mu), i)
        cin.set(c0)


@pytest.mark.parametrize("cnt,shared", [(3, False), (3, True), (600, False)])
def test_cmux_ntt(engine, cnt, shared):
    """TRGSW2NTT + CMUXNTT (res = c0 + (c1 - c0) [.] cs), bytes against the
    oracle's exact product; per-item and shared TRGSW; both kernel instantiations; extreme
    operands; out aliasing c0."""
    cf = engine
    rng = np.random.default_rng(11 + cnt)
    ncs = 1 if shared else cnt
    tg = rng.integers(0, 2**32, (ncs, 4, 2, 1024), dtype=np.uint64).astype(np.uint32)
    c1 = rng.integers(0, 2**32, (cnt, 2048), dtype=np.uint64).astype(np.uint32)
    c0 = rng.integers(0, 2**32, (cnt, 2048), dtype=np.uint64).astype(np.uint32)
    tg[0, 1] = 0xFFFFFFFF                   # maximal key coefficients in one row
    c1[1], c0[1] = 0x7FFFFFFF, 0x80000000   # every digit extreme
    c1[2] = c0[2]                           # c1 - c0 = 0: digits = rounding only
    st = cf.Stream(0).Create()
    cs = cf.cuFHETRGSWNTTlvl1(ncs)
    cf.TRGSW2NTT(cs, tg, st)
    t1, t0, res = cf.cuFHETRLWElvl1(cnt).set(c1), cf.cuFHETRLWElvl1(cnt).set(c0), cf.cuFHETRLWElvl1(cnt)
    cf.CMUXNTT(res, cs, t1, t0, st)
    cf.gCMUXNTT(t0, cs, t1, t0, st)         # in place: out aliases c0
    with torch.cuda.stream(st.torch_stream()):
        t0._host_t.copy_(t0.trlwedevice, non_blocking=True)
    cf.Synchronize()
    for i in (range(cnt) if cnt < 10 else (0, 1, 2, 3, 299, 598, 599)):
        want = orc.cmux(tg[0 if shared else i], c1[i], c0[i])
        assert np.array_equal(res.trlwehost[i], want), i
        assert np.array_equal(t0.trlwehost[i], want), ("alias", i)


# ------------------------------------------------------------------- gates ---
@pytest.mark.parametrize("level", [0, 1])
def test_all_gates_bit_exact_and_decrypt(engine, keys, level):
    """13 gates + NMux (test/test_gate_gpu.cc:36-91, truth tables test/plain.h:10-69)."""
    cf = engine
    cnt = 12
    rng = np.random.default_rng(100 + level)
    for op_name in orc.OPS:
        op = orc.OP[op_name]
        b0, b1, b2 = (rng.integers(0, 2, cnt) for _ in range(3))
        x0 = orc.encrypt(keys, b0, level, base_index=1000 * op)
        x1 = orc.encrypt(keys, b1, level, base_index=1000 * op + 300)
        x2 = orc.encrypt(keys, b2, level, base_index=1000 * op + 600)
        unary = op_name in ("NOT", "COPY")
        ternary = op_name in ("MUX", "NMUX")
        ins = [x0, None if unary else x1, x2 if ternary else None]
        got = run_gate(cf, op_name, level, ins, cnt)
        ref = orc.gate_batch(keys, op, level, x0, None if unary else x1, x2 if ternary else None)
        assert np.array_equal(got, ref), f"{op_name} level {level}: ciphertext bytes differ"
        dec = orc.decrypt(keys, got, level)
        truth = [orc.truth(op, b0[i], b1[i], b2[i]) for i in range(cnt)]
        assert list(dec) == truth, f"{op_name} level {level}: wrong plaintext"


@pytest.mark.parametrize("level", [0, 1])
def test_mixed_batch(engine, keys, level):
    """BASELINE config 4 shape: op[i] = i mod 14 over one batch."""
    cf = engine
    cnt = 45
    rng = np.random.default_rng(7 + level)
    ops = (np.arange(cnt) % 14).astype(np.uint8)
    b = [rng.integers(0, 2, cnt) for _ in range(3)]
    x = [orc.encrypt(keys, b[k], level, base_index=50000 + 1000 * k) for k in range(3)]
    st = cf.Stream(0).Create()
    cts = [cf.Ctxt(cnt, level).set(v) for v in x]
    out = cf.Ctxt(cnt, level)
    for c in cts:
        cf.CtxtCopyH2D(c, st)
    cf.gGateMixed(ops, out, cts[0], cts[1], cts[2], st)
    cf.CtxtCopyD2H(out, st)
    cf.Synchronize()
    ref = orc.gate_batch(keys, 0, level, x[0], x[1], x[2], ops=ops)
    assert np.array_equal(out.tlwehost, ref)
    dec = orc.decrypt(keys, out.tlwehost, level)
    assert list(dec) == [orc.truth(ops[i], b[0][i], b[1][i], b[2][i]) for i in range(cnt)]


@pytest.mark.parametrize("level,cnt", [(0, 14 * 45), (1, 14 * 64)])
def test_mixed_batch_across_waves(engine, keys, level, cnt):
    """A mixed batch whose blind-rotation work items (cnt of them: Mux/NMux count twice, Not/Copy not
    at all) fill one wave of 4 gates per SM on a B200 plus a partial one, so the per-item lists go
    through the separate launch of the last partial wave (one-gate CTAs for 630, three-gate CTAs
    for 896) and the in-step key switch: every gate decrypts right, a strided sample matches the
    oracle's bytes."""
    cf = engine
    rng = np.random.default_rng(17 + level)
    ops = (np.arange(cnt) % 14).astype(np.uint8)
    b = [rng.integers(0, 2, cnt) for _ in range(3)]
    x = [orc.encrypt(keys, b[k], level, base_index=70000 + 2000 * k) for k in range(3)]
    st = cf.Stream(0).Create()
    cts = [cf.Ctxt(cnt, level).set(v) for v in x]
    out = cf.Ctxt(cnt, level)
    for c in cts:
        cf.CtxtCopyH2D(c, st)
    cf.gGateMixed(ops, out, cts[0], cts[1], cts[2], st)
    cf.CtxtCopyD2H(out, st)
    cf.Synchronize()
    dec = orc.decrypt(keys, out.tlwehost, level)
    assert list(dec) == [orc.truth(ops[i], b[0][i], b[1][i], b[2][i]) for i in range(cnt)]
    idx = np.concatenate([np.arange(0, cnt, 23), np.arange(cnt - 14, cnt)])
    ref = orc.gate_batch(keys, 0, level, x[0][idx], x[1][idx], x[2][idx], ops=ops[idx])
    assert np.array_equal(out.tlwehost[idx], ref)


def test_aliasing_and_stream_chain(engine, keys):
    """out may alias an input; gates on one stream run in issue order
    (test/test_intensive.cc:21-54: 20 back-to-back NANDs per stream)."""
    cf = engine
    cnt = 6
    st = cf.Stream(0).Create()
    a = cf.Ctxt(cnt, 0).set(orc.encrypt(keys, [1] * cnt, 0, base_index=9000))
    b = cf.Ctxt(cnt, 0).set(orc.encrypt(keys, [0, 1] * (cnt // 2), 0, base_index=9100))
    cf.CtxtCopyH2D(a, st)
    cf.CtxtCopyH2D(b, st)
    ref = a.tlwehost.copy()
    for _ in range(5):
        cf.gNand(a, a, b, st)                  # a = NAND(a, b) in place
        ref = orc.gate_batch(keys, orc.OP["NAND"], 0, ref, b.tlwehost)
    cf.CtxtCopyD2H(a, st)
    while not cf.StreamQuery(st):
        pass
    assert np.array_equal(a.tlwehost, ref)


def test_host_buffer_entry_point(engine, keys):
    """b2fhe_gate_batch_host = CtxtCopyH2D + gate + CtxtCopyD2H (src/cufhe_gates_gpu.cu:148-158)."""
    cf = engine
    L = cf._lib.lib()
    cnt = 9
    x0 = orc.encrypt(keys, [0, 1, 1] * 3, 0, base_index=777)
    x1 = orc.encrypt(keys, [1, 1, 0] * 3, 0, base_index=888)
    out = np.zeros_like(x0)
    cf._lib.check(L.b2fhe_gate_batch_host(0, None, orc.OP["XOR"], None, 0, out.ctypes.data, x0.ctypes.data,
                                          x1.ctypes.data, None, cnt))
    cf.Synchronize()
    assert np.array_equal(out, orc.gate_batch(keys, orc.OP["XOR"], 0, x0, x1))


def test_error_paths(engine):
    cf = engine
    L = cf._lib.lib()
    assert L.b2fhe_gate_batch(0, None, 99, 0, C.c_void_p(8), C.c_void_p(8), None, None, 1) != 0
    assert b"b2fhe" in L.b2fhe_last_error_string()
    assert L.b2fhe_gate_batch(0, None, 0, 0, C.c_void_p(8), C.c_void_p(8), None, None, 1) != 0   # NAND without in1
    assert L.b2fhe_gate_batch(5, None, 0, 0, None, None, None, None, 1) != 0                      # device not initialised


# ------------------------------------------------ full-size property tests ---
def test_full_batch_16384_nand(engine, keys):
    """BASELINE config 3: 16384 independent NANDs.  100 % decryption agreement over
    the whole batch, byte parity with the oracle on a strided sample, and
    batch-position independence (same inputs at two positions -> same bytes)."""
    cf = engine
    cnt = 16384
    rng = np.random.default_rng(3)
    b0, b1 = rng.integers(0, 2, cnt), rng.integers(0, 2, cnt)
    x0 = orc.encrypt(keys, b0, 0, base_index=200000)
    x1 = orc.encrypt(keys, b1, 0, base_index=300000)
    x0[-1], x1[-1] = x0[0], x1[0]
    b0[-1], b1[-1] = b0[0], b1[0]
    got = run_gate(cf, "NAND", 0, [x0, x1, None], cnt)
    assert np.array_equal(orc.decrypt(keys, got, 0), 1 - b0 * b1)
    assert np.array_equal(got[-1], got[0])
    idx = np.arange(0, cnt, 431)
    ref = orc.gate_batch(keys, orc.OP["NAND"], 0, x0[idx], x1[idx])
    assert np.array_equal(got[idx], ref)


def test_cluster_kernel_stress_against_throughput_kernel(engine):
    """The cluster kernel hands data between CTAs with 128-bit st.async stores / mbarriers and
    reuses double-buffered receive areas every other step: a protocol slip would show up as sporadic
    corruption.  One CTA per SM (33 clusters), two per SM (70 clusters) and repeated launches on random
    inputs must reproduce the throughput kernel's bits every time."""
    cf = engine
    L = cf._lib.lib()
    rng = np.random.default_rng(77)
    try:
        for cnt in (33, 70, 7):
            x = torch.from_numpy(rng.integers(0, 2**16, (cnt, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
            want = torch.empty((cnt, 1025), dtype=torch.int32, device="cuda")
            cf._lib.check(L.b2fhe_set_br_variant(40))
            cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, want.data_ptr(), x.data_ptr(), 1 << 29, cnt))
            cf._lib.check(L.b2fhe_set_br_variant(5))
            for rep in range(6):
                got = torch.zeros_like(want)
                cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, got.data_ptr(), x.data_ptr(), 1 << 29, cnt))
                torch.cuda.synchronize()
                assert torch.equal(got, want), (cnt, rep)
    finally:
        cf._lib.check(L.b2fhe_set_br_variant(0))


@pytest.mark.gpu
def test_default_large_batch_paths_against_single_row_kernel(engine):
    """The default large-batch path = k_blind_rotate_s<4> in full waves + a separate launch (fewer gates
    per CTA) for a last partial wave, then the key switch.  Batch sizes around the wave boundaries (one
    gate over, one SM over, just below the no-split threshold, idle warps in the last key-switch CTA)
    must reproduce the bits of the one-gate-per-CTA launch and of k_keyswitch, with and without the
    split (the oracle pins the kernels themselves in test_blind_rotate_and_keyswitch_stages)."""
    cf = engine
    L = cf._lib.lib()
    rng = np.random.default_rng(99)
    sm = torch.cuda.get_device_properties(0).multi_processor_count
    wave = 4 * sm
    try:
        for cnt in (wave + 1, wave + sm + 1, wave + 3 * sm, wave + 3 * sm + 1, 2 * wave, 2 * wave - 5):
            x = torch.from_numpy(rng.integers(0, 2**16, (cnt, 637), dtype=np.uint64).astype(np.uint16).view(np.int16)).cuda()
            want = torch.empty((cnt, 1025), dtype=torch.int32, device="cuda")
            wks = torch.empty((cnt, 637), dtype=torch.int16, device="cuda")
            cf._lib.check(L.b2fhe_set_br_variant(40))
            cf._lib.check(L.b2fhe_set_gates_per_cta(1))              # same kernel, one gate per CTA, one launch
            cf._lib.check(L.b2fhe_set_ks_variant(2))
            cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, want.data_ptr(), x.data_ptr(), 1 << 29, cnt))
            cf._lib.check(L.b2fhe_keyswitch_batch(0, None, wks.data_ptr(), want.data_ptr(), cnt))
            cf._lib.check(L.b2fhe_set_br_variant(0))
            cf._lib.check(L.b2fhe_set_gates_per_cta(0))
            cf._lib.check(L.b2fhe_set_ks_variant(0))
            for split in (-2, -1):
                cf._lib.check(L.b2fhe_set_br_variant(split))
                got = torch.zeros_like(want)
                gks = torch.zeros_like(wks)
                cf._lib.check(L.b2fhe_blind_rotate_batch(0, None, got.data_ptr(), x.data_ptr(), 1 << 29, cnt))
                cf._lib.check(L.b2fhe_keyswitch_batch(0, None, gks.data_ptr(), got.data_ptr(), cnt))
                torch.cuda.synchronize()
                assert torch.equal(got, want), (cnt, split)
                assert torch.equal(gks, wks), (cnt, split)
    finally:
        cf._lib.check(L.b2fhe_set_br_variant(-2))
        cf._lib.check(L.b2fhe_set_br_variant(0))
        cf._lib.check(L.b2fhe_set_gates_per_cta(0))
        cf._lib.check(L.b2fhe_set_ks_variant(0))


@pytest.mark.gpu
def test_default_keyswitch_step_above_latency_threshold(engine, keys):
    """Above 32 gates per SM the default key switch is k_keyswitch_step<16, 2> (below: k_keyswitch_lat).
    A batch just over the threshold with a PARTIAL last CTA: every gate compared with the one-warp-per-gate
    k_keyswitch (variant 2), the first ones and the extreme rows with the oracle."""
    cf = engine
    L = cf._lib.lib()
    sm = torch.cuda.get_device_properties(0).multi_processor_count
    cnt = 32 * sm + 37                      # 4773 on a B200: 298 full CTAs of 16 gates + one of 5
    rng = np.random.default_rng(123)
    x = rng.integers(0, 2**32, (cnt, 1025), dtype=np.uint64).astype(np.uint32)
    x[1, :] = 0
    x[2, :] = 0xFFFFFFFF
    x[cnt - 1, :] = 0x80000000
    dx = dev_u32(x)
    want = torch.zeros((cnt, 637), dtype=torch.int16, device="cuda")
    got = torch.zeros_like(want)
    try:
        cf._lib.check(L.b2fhe_set_ks_variant(2))
        cf._lib.check(L.b2fhe_keyswitch_batch(0, None, want.data_ptr(), dx.data_ptr(), cnt))
        for variant in (0, 21):
            cf._lib.check(L.b2fhe_set_ks_variant(variant))
            got.zero_()
            cf._lib.check(L.b2fhe_keyswitch_batch(0, None, got.data_ptr(), dx.data_ptr(), cnt))
            torch.cuda.synchronize()
            assert torch.equal(got, want), variant
        # k_keyswitch_lat on both sides of ITS 32-warp / 16-warp switch (4 gates per SM), default and round-1 form
        for variant, n in ((0, 4 * sm), (0, 4 * sm + 3), (1, 4 * sm + 3), (3, 4 * sm + 3), (3, 9)):
            cf._lib.check(L.b2fhe_set_ks_variant(variant))
            got.zero_()
            cf._lib.check(L.b2fhe_keyswitch_batch(0, None, got.data_ptr(), dx.data_ptr(), n))
            torch.cuda.synchronize()
            assert torch.equal(got[:n], want[:n]), (variant, n)
        cf._lib.check(L.b2fhe_set_ks_variant(0))
        got.zero_()
        cf._lib.check(L.b2fhe_keyswitch_batch(0, None, got.data_ptr(), dx.data_ptr(), cnt))
        torch.cuda.synchronize()
    finally:
        cf._lib.check(L.b2fhe_set_ks_variant(0))
    g = to_np(got, np.uint16)
    for i in (0, 1, 2, 3, cnt - 1):
        assert np.array_equal(g[i], orc.keyswitch(keys, x[i])), i


@pytest.mark.gpu
@pytest.mark.parametrize("level", [0, 1])
def test_mixed_batch_many_waves_equals_uniform_paths(engine, keys, level):
    """BASELINE config 4 at a size where the host-built work lists cross SEVERAL waves of the blind rotation
    (5 full waves + a partial one; Mux / NMux add two work items each): every gate of the mixed batch must
    equal the same gate evaluated by the uniform per-op path (which the oracle and the reference GPU build pin
    byte for byte on smaller batches), and decrypt to the truth table."""
    cf = engine
    L = cf._lib.lib()
    sm = torch.cuda.get_device_properties(0).multi_processor_count
    cnt = 5 * 4 * sm + 77
    rng = np.random.default_rng(11 + level)
    ops = (np.arange(cnt) % 14).astype(np.uint8)
    b = [rng.integers(0, 2, cnt) for _ in range(3)]
    x = [orc.encrypt(keys, b[k], level, base_index=2000000 + 100000 * k) for k in range(3)]
    tdt = torch.int16 if level == 0 else torch.int32
    dt = np.int16 if level == 0 else np.int32
    d = [torch.from_numpy(v.view(dt)).cuda() for v in x]
    out = torch.zeros_like(d[0])
    cf._lib.check(L.b2fhe_gate_batch_mixed(0, None, ops.ctypes.data, level, out.data_ptr(), d[0].data_ptr(), d[1].data_ptr(),
                                           d[2].data_ptr(), cnt))
    torch.cuda.synchronize()
    got = out.cpu().numpy().view(x[0].dtype)
    for op in range(14):
        idx = np.nonzero(ops == op)[0]
        sub = [torch.from_numpy(np.ascontiguousarray(v[idx]).view(dt)).cuda() for v in x]
        o = torch.zeros((len(idx), x[0].shape[1]), dtype=tdt, device="cuda")
        unary, ternary = op >= 12, op in (10, 11)
        cf._lib.check(L.b2fhe_gate_batch(0, None, op, level, o.data_ptr(), sub[0].data_ptr(), None if unary else sub[1].data_ptr(),
                                         sub[2].data_ptr() if ternary else None, len(idx)))
        torch.cuda.synchronize()
        assert np.array_equal(got[idx], o.cpu().numpy().view(x[0].dtype)), f"op {op} level {level}"
    dec = orc.decrypt(keys, got, level)
    assert list(dec) == [orc.truth(int(ops[i]), int(b[0][i]), int(b[1][i]), int(b[2][i])) for i in range(cnt)]
