"""Pins the oracle against the REAL reference CPU library: keys and ciphertexts made by
TFHEpp itself (oracle/_ref, built from /root/reference), gates evaluated by the oracle,
results decrypted by TFHEpp.  SURVEY.md F4: TFHEpp's ciphertext bits differ from cuFHE's
by design (rounding offset, FFT), so the pin is at decryption level for every gate and
both I/O levels, plus agreement with TFHEpp's own HomNAND."""
import numpy as np
import pytest

from oracle import orc, tfhepp

pytestmark = pytest.mark.skipif(not tfhepp.available(), reason="oracle/_ref not built (needs /root/reference)")


@pytest.fixture(scope="module")
def tkeys():
    return tfhepp.Keys()


def test_key_layout_statistics(tkeys):
    """TFHEpp's exported raw arrays have the layout the engine assumes: binary lvl0 key,
    ternary lvl1 key, and key-switching-key rows that decrypt (as lvl0 TLWEs) to
    s1[j] * (v+1) * 2^(16 - 2(d+1)) (TFHEpp/include/evalkeygens.hpp:188-203)."""
    assert set(np.unique(tkeys.sk0)) <= {0, 1} and set(np.unique(tkeys.sk1)) <= {-1, 0, 1}
    ksk = tkeys.ksk.reshape(1024, 7, 2, 637)
    for j in (0, 5, 1023):
        for d in (0, 3, 6):
            for v in (0, 1):
                row = ksk[j, d, v].astype(np.int64)
                phase = (row[636] - row[:636] @ tkeys.sk0.astype(np.int64)) & 0xFFFF
                want = (int(tkeys.sk1[j]) * (v + 1) * (1 << (16 - 2 * (d + 1)))) & 0xFFFF
                err = ((phase - want + 0x8000) & 0xFFFF) - 0x8000
                assert abs(err) < 64, (j, d, v, err)


@pytest.mark.parametrize("level", [0, 1])
def test_oracle_gates_on_tfhepp_keys_decrypt_by_tfhepp(tkeys, level):
    rng = np.random.default_rng(level)
    cnt = 8
    for name in orc.OPS:
        op = orc.OP[name]
        b = [rng.integers(0, 2, cnt).astype(np.uint8) for _ in range(3)]
        x = [tfhepp.encrypt(b[k], level) for k in range(3)]
        out = orc.gate_batch(tkeys, op, level, x[0], x[1], x[2])
        dec = tfhepp.decrypt(out, level)
        assert list(dec) == [orc.truth(op, b[0][i], b[1][i], b[2][i]) for i in range(cnt)], name


def test_oracle_agrees_with_tfhepp_homnand_at_decrypt_level(tkeys):
    rng = np.random.default_rng(5)
    b0, b1 = rng.integers(0, 2, 16).astype(np.uint8), rng.integers(0, 2, 16).astype(np.uint8)
    x0, x1 = tfhepp.encrypt(b0, 0), tfhepp.encrypt(b1, 0)
    ours = tfhepp.decrypt(orc.gate_batch(tkeys, orc.OP["NAND"], 0, x0, x1), 0)
    theirs = tfhepp.decrypt(tfhepp.homnand_lvl0(x0, x1), 0)
    assert list(ours) == list(theirs) == list(1 - b0 * b1)


# ---- wire formats: byte-for-byte against TFHEpp's own cereal output ----
def _wire():
    import ctypes as C
    from cufhe_b200 import _lib
    _lib.build()
    return _lib, _lib.lib(), C


def test_wire_evalkey_reader_and_writer_match_tfhepp(tkeys):
    _lib, L, C = _wire()
    ref = tfhepp.serialize_evalkey(0)                 # bk<lvl01> + iksk<lvl10>, what cuFHE's tests build
    assert L.b2fhe_wire_evalkey_size() == ref.size
    # reader: finds TFHEpp's raw arrays inside TFHEpp's archive
    ob, ok = C.c_size_t(), C.c_size_t()
    _lib.check(L.b2fhe_wire_evalkey_parse(ref.ctypes.data, ref.size, C.byref(ob), C.byref(ok)))
    assert np.array_equal(ref[ob.value:ob.value + tkeys.bk.nbytes], tkeys.bk.view(np.uint8))
    assert np.array_equal(ref[ok.value:ok.value + tkeys.ksk.nbytes], tkeys.ksk.view(np.uint8))
    # writer: same bytes as TFHEpp's writer (but for the 8 bytes of lvl1.alpha, which TFHEpp never
    # initialises -- lweParams.hpp:43-50)
    mine = np.zeros(ref.size, np.uint8)
    wrote = C.c_size_t()
    _lib.check(L.b2fhe_wire_evalkey_write(mine.ctypes.data, mine.size, tkeys.bk.ctypes.data, tkeys.ksk.ctypes.data, C.byref(wrote)))
    assert wrote.value == ref.size
    a, b = mine.copy(), ref.copy()
    a[33:41] = 0; b[33:41] = 0
    assert np.array_equal(a, b)
    # and TFHEpp's own reader loads what the writer wrote
    rc, bk2, ksk2 = tfhepp.deserialize_evalkey(mine)
    assert rc == 0 and np.array_equal(bk2, tkeys.bk) and np.array_equal(ksk2, tkeys.ksk)


def test_wire_evalkey_reader_skips_fft_key_and_rejects_garbage(tkeys):
    _lib, L, C = _wire()
    big = tfhepp.serialize_evalkey(1)                 # TFHEpp tutorial's cloud.key also carries bkfft
    ob, ok = C.c_size_t(), C.c_size_t()
    _lib.check(L.b2fhe_wire_evalkey_parse(big.ctypes.data, big.size, C.byref(ob), C.byref(ok)))
    assert np.array_equal(big[ob.value:ob.value + tkeys.bk.nbytes], tkeys.bk.view(np.uint8))
    assert np.array_equal(big[ok.value:ok.value + tkeys.ksk.nbytes], tkeys.ksk.view(np.uint8))
    small = big[:4096].copy()
    assert L.b2fhe_wire_evalkey_parse(small.ctypes.data, small.size, C.byref(ob), C.byref(ok)) == _lib.E_BAD_FORMAT
    bad = big[:200].copy(); bad[1] ^= 0xFF            # lvl0.n
    assert L.b2fhe_wire_evalkey_parse(bad.ctypes.data, bad.size, C.byref(ob), C.byref(ok)) == _lib.E_BAD_FORMAT
    assert b"parameter set" in L.b2fhe_last_error_string()
    be = big[:200].copy(); be[0] = 0
    assert L.b2fhe_wire_evalkey_parse(be.ctypes.data, be.size, C.byref(ob), C.byref(ok)) == _lib.E_BAD_FORMAT


@pytest.mark.parametrize("level", [0, 1])
def test_wire_ciphertext_vectors_match_tfhepp(tkeys, level):
    _lib, L, C = _wire()
    for cnt in (0, 1, 5):
        cts = tfhepp.encrypt(np.arange(cnt) % 2, level) if cnt else (
            np.zeros((0, 637), np.uint16) if level == 0 else np.zeros((0, 1025), np.uint32))
        ref = tfhepp.serialize_tlwes(cts, level)
        assert L.b2fhe_wire_tlwes_size(level, cnt) == ref.size
        mine = np.zeros(ref.size, np.uint8)
        _lib.check(L.b2fhe_wire_tlwes_write(mine.ctypes.data, mine.size, level, cts.ctypes.data, cnt, None))
        assert np.array_equal(mine, ref)
        n, off = C.c_size_t(), C.c_size_t()
        _lib.check(L.b2fhe_wire_tlwes_parse(ref.ctypes.data, ref.size, level, C.byref(n), C.byref(off)))
        assert n.value == cnt and np.array_equal(ref[off.value:], cts.view(np.uint8).reshape(-1))
        got_n, got = tfhepp.deserialize_tlwes(mine, level, max(cnt, 1))
        assert got_n == cnt and np.array_equal(got[:cnt], cts)
        if cnt:      # a level-0 archive is not a level-1 archive
            assert L.b2fhe_wire_tlwes_parse(ref.ctypes.data, ref.size, 1 - level, C.byref(n), C.byref(off)) == _lib.E_BAD_FORMAT
