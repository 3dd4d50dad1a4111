"""Levelised circuits on device-resident batches (BASELINE config 5):
ripple-carry adder over several instances -- decrypts to a + b, and every output
ciphertext is byte-identical to the same gate sequence evaluated by the CPU oracle."""
import numpy as np
import pytest

from oracle import orc

pytestmark = pytest.mark.gpu


def oracle_adder(keys, level, A, B):
    """Same netlist as cufhe_b200.circuits.ripple_carry_add, gate by gate on the oracle.
    A, B: [n][inst][words]."""
    n = A.shape[0]
    G = lambda op, x, y: orc.gate_batch(keys, orc.OP[op], level, x, y)   # noqa: E731
    x = [G("XOR", A[i], B[i]) for i in range(n)]
    g = [G("AND", A[i], B[i]) for i in range(n - 1)]
    c = [None, g[0]]
    for i in range(1, n - 1):
        t = G("AND", c[i], x[i])
        c.append(G("OR", g[i], t))
    return np.stack([x[0]] + [G("XOR", x[i], c[i]) for i in range(1, n)])


@pytest.mark.parametrize("level,nbits,inst", [(0, 8, 5), (1, 4, 3)])
def test_ripple_carry_adder(engine, keys, level, nbits, inst):
    cf = engine
    from cufhe_b200 import circuits
    rng = np.random.default_rng(7 + level)
    av = rng.integers(0, 2**nbits, inst)
    bv = rng.integers(0, 2**nbits, inst)
    av[0], bv[0] = 2**nbits - 1, 1                    # full carry propagation + wrap-around
    abits = np.array([[(int(v) >> i) & 1 for v in av] for i in range(nbits)], np.uint8)
    bbits = np.array([[(int(v) >> i) & 1 for v in bv] for i in range(nbits)], np.uint8)
    A = np.stack([orc.encrypt(keys, abits[i], level, base_index=1000 * i) for i in range(nbits)])
    B = np.stack([orc.encrypt(keys, bbits[i], level, base_index=1000 * i + 500) for i in range(nbits)])
    st = cf.Stream(0).Create()
    wa = circuits.Wires(nbits, inst, level).load(A)
    wb = circuits.Wires(nbits, inst, level).load(B)
    s, gates = circuits.ripple_carry_add(wa, wb, st)
    cf.Synchronize()
    assert gates == circuits.adder_gate_count(nbits) * inst
    got = s.numpy()
    dec = np.stack([orc.decrypt(keys, got[i], level) for i in range(nbits)])
    sums = [sum(int(dec[i, k]) << i for i in range(nbits)) for k in range(inst)]
    assert sums == [int((a + b) % 2**nbits) for a, b in zip(av, bv)]
    assert np.array_equal(got, oracle_adder(keys, level, A, B))


def test_32_bit_adder_many_instances_decrypts(engine, keys):
    """BASELINE config 5 at full width: 32-bit ripple-carry additions over 96 instances (157 bootstraps each,
    65 dependent levels of batched launches), checked by decryption against a + b mod 2^32 -- the byte check
    against the oracle is the small case above (the oracle needs seconds per gate batch)."""
    cf = engine
    from cufhe_b200 import circuits
    nbits, inst, level = 32, 96, 0
    rng = np.random.default_rng(32)
    av = rng.integers(0, 2**32, inst, dtype=np.uint64)
    bv = rng.integers(0, 2**32, inst, dtype=np.uint64)
    av[0], bv[0] = 2**32 - 1, 1                       # carry through all 32 bits
    av[1], bv[1] = 0, 0
    abits = np.array([[(int(v) >> i) & 1 for v in av] for i in range(nbits)], np.uint8)
    bbits = np.array([[(int(v) >> i) & 1 for v in bv] for i in range(nbits)], np.uint8)
    A = np.stack([orc.encrypt(keys, abits[i], level, base_index=7000000 + 1000 * i) for i in range(nbits)])
    B = np.stack([orc.encrypt(keys, bbits[i], level, base_index=8000000 + 1000 * i) for i in range(nbits)])
    st = cf.Stream(0).Create()
    wa = circuits.Wires(nbits, inst, level).load(A)
    wb = circuits.Wires(nbits, inst, level).load(B)
    s, gates = circuits.ripple_carry_add(wa, wb, st)
    cf.Synchronize()
    assert gates == circuits.adder_gate_count(nbits) * inst
    got = s.numpy()
    dec = np.stack([orc.decrypt(keys, got[i], level) for i in range(nbits)])
    sums = [sum(int(dec[i, k]) << i for i in range(nbits)) for k in range(inst)]
    assert sums == [int((int(a) + int(b)) % 2**32) for a, b in zip(av, bv)]
