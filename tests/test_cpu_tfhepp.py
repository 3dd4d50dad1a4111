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
