"""CPU-side tests (no GPU): the oracle against the reference's own truths, the
warp-level emulation of the CUDA arithmetic, host logic, and the C-ABI surface."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N = 1024


# --------------------------------------------------------------- the oracle ---
def test_oracle_polymul_ntt_matches_schoolbook():
    """The reference's only arithmetic KAT: NTT product == schoolbook negacyclic product
    mod 2^32 (test/test_polynomial_mult_1024.cu:51-73,209-223; we demand diff 0)."""
    rng = np.random.default_rng(0)
    for t in range(6):
        a = rng.integers(0, 2**32, N, dtype=np.uint64).astype(np.uint32)
        hi = 256 if t < 3 else 2**18          # the reference's 8-bit and "TFHE-safe" 18-bit shapes
        b = rng.integers(0, hi, N).astype(np.uint32)
        assert np.array_equal(orc.polymul(a, b, True), orc.polymul(a, b, False))


def test_oracle_polymul_known_answers():
    a = np.zeros(N, np.uint32); b = np.zeros(N, np.uint32)
    a[1] = 5; b[N - 1] = 3                      # 5X * 3X^1023 = 15 X^1024 = -15
    out = orc.polymul(a, b, True)
    assert out[0] == np.uint32(-15 & 0xFFFFFFFF) and not out[1:].any()
    a[:] = 0; a[0] = 0xFFFFFFFF; b[:] = 0; b[0] = 2   # wrap mod 2^32
    assert orc.polymul(a, b, False)[0] == 0xFFFFFFFE


def test_oracle_cmux_ntt_matches_schoolbook(keys):
    rng = np.random.default_rng(1)
    acc = rng.integers(0, 2**32, (2, N), dtype=np.uint64).astype(np.uint32)
    bk_i = keys.bk[: 8 * N]
    for abar in (0, 1, 1023, 1024, 1500, 2047):
        assert np.array_equal(orc.cmux_step(acc, abar, bk_i, True), orc.cmux_step(acc, abar, bk_i, False))
    assert np.array_equal(orc.cmux_step(acc, 0, bk_i, False), acc)   # abar = 0 is a no-op


@pytest.mark.parametrize("level", [0, 1])
def test_oracle_gates_decrypt_to_truth_tables(keys, level):
    """test/plain.h:10-69 through decryption, all 14 ops, all input combinations."""
    for name in orc.OPS:
        op = orc.OP[name]
        combos = [(a, b, c) for a in (0, 1) for b in (0, 1) for c in (0, 1)]
        x = [orc.encrypt(keys, [cb[k] for cb in combos], level, base_index=100 * op + 10 * k) for k in range(3)]
        out = orc.gate_batch(keys, op, level, x[0], x[1], x[2])
        dec = orc.decrypt(keys, out, level)
        assert list(dec) == [orc.truth(op, *cb) for cb in combos], name


def test_oracle_schoolbook_blind_rotate_equals_ntt(keys):
    c = orc.encrypt(keys, [1], 0, base_index=5)[0]
    c[40:] = 0     # keep the O(N^2) path short: abar = 0 steps are no-ops either way
    assert np.array_equal(orc.blind_rotate_extract(keys, c, schoolbook=True),
                          orc.blind_rotate_extract(keys, c, schoolbook=False))


def test_golden_fixture(keys):
    """Committed fixture (tests/golden/make_golden.py): digests of oracle outputs on the
    seeded key set -- guards the oracle and the key generator against drift."""
    import hashlib, json
    with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as f:
        G = json.load(f)
    assert hashlib.sha256(keys.bk.tobytes()).hexdigest() == G["bk_sha256"]
    assert hashlib.sha256(keys.ksk.tobytes()).hexdigest() == G["ksk_sha256"]
    for level in (0, 1):
        x0 = orc.encrypt(keys, [0, 1, 0, 1], level, base_index=1)
        x1 = orc.encrypt(keys, [0, 0, 1, 1], level, base_index=101)
        out = orc.gate_batch(keys, orc.OP["NAND"], level, x0, x1)
        assert hashlib.sha256(out.tobytes()).hexdigest() == G[f"nand_lvl{level}_sha256"]
        assert out[0, :4].tolist() == G[f"nand_lvl{level}_first_words"]


def test_oracle_trlwe_level_and_cmux(keys):
    """SURVEY.md 8f ranks 2-3 on the oracle: TLWE2TRLWE / SampleExtract+KeySwitch / Refresh keep
    the plaintext; CMUX by exact NTT == schoolbook, and selects c1 / c0 at decrypt level."""
    c = orc.encrypt(keys, [1, 0], 0, base_index=31)
    for i, bit in enumerate((1, 0)):
        t = orc.blind_rotate_trlwe(keys, c[i])
        e = orc.blind_rotate_extract(keys, c[i])
        assert e[0] == t[0] and e[N] == t[N] and e[1] == (-int(t[N - 1])) % 2**32
        back = orc.sample_extract_keyswitch(keys, orc.refresh(keys, t))
        assert orc.decrypt(keys, back[None], 0)[0] == bit
    rng = np.random.default_rng(5)
    tg = rng.integers(0, 2**32, (4, 2, N), dtype=np.uint64).astype(np.uint32)
    c1 = rng.integers(0, 2**32, 2 * N, dtype=np.uint64).astype(np.uint32)
    c0 = rng.integers(0, 2**32, 2 * N, dtype=np.uint64).astype(np.uint32)
    assert np.array_equal(orc.cmux(tg, c1, c0, schoolbook=False), orc.cmux(tg, c1, c0, schoolbook=True))
    # selector semantics with the real key: BK_i encrypts key bit s_i, so CMUX(BK_i, c1, c0)
    # decrypts like c1 when s_i = 1 and like c0 when s_i = 0
    t1 = orc.blind_rotate_trlwe(keys, c[0])          # encrypts +mu pattern (bit 1)
    t0 = orc.blind_rotate_trlwe(keys, c[1])          # bit 0
    bk = keys.bk.reshape(636, 4, 2, N)
    for i in (0, 1, 2, 3):
        sel = orc.cmux(bk[i], t1, t0)
        got = orc.decrypt(keys, orc.sample_extract_keyswitch(keys, sel)[None], 0)[0]
        assert got == (1 if keys.sk0[i] else 0), i


def test_adder_netlist_gate_count():
    from cufhe_b200.circuits import adder_gate_count
    assert adder_gate_count(32) == 32 + 31 + 2 * 30 + 31
    assert adder_gate_count(8) == 8 + 7 + 12 + 7


# ------------------------------------- CUDA arithmetic, emulated on the CPU ---
def test_team_emulation_of_signed_ntt_cmux_and_exact_product(tmp_path):
    """Compiles the SAME host+device headers the throughput kernels use (csrc/ntt16s.cuh: signed lazy
    residues, 128-bit exchanges) and runs the 64-thread team dataflow of k_blind_rotate_s /
    k_polymul_s thread by thread against the oracle: CMUX steps, two-prime and exact three-prime
    products, bank-conflict freedom of every exchange, interval bounds of the lazy residues."""
    orc.build()
    exe = str(tmp_path / "emul_ntt16s")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "emul", "emul_ntt16s.cpp"),
                           "-L" + os.path.join(ROOT, "oracle"), "-lorc", "-Wl,-rpath," + os.path.join(ROOT, "oracle"),
                           "-fopenmp"])
    r = subprocess.run([exe, "6"], capture_output=True, text=True)
    out = r.stdout
    assert r.returncode == 0, out
    assert "primitive mismatches (SOL == IMAD form, residues, ranges): 0" in out
    assert "polymul mismatches: 0 / 6" in out and "three-prime exact polymul mismatches: 0 / 6" in out
    assert "cmux mismatches: 0 / 6" in out and "bank conflicts: 0" in out


def test_step_schedule_race_detector_on_the_real_kernel_source(tmp_path):
    """tests/emul/emul_sched.cpp compiles the REAL CMUX step of k_blind_rotate_s (brs_gate_loop and everything
    below it, B2_HOST_TRACE build of the same headers) for the host and runs it with one host thread per CUDA
    thread -- every lane independent -- under a vector-clock race detector over the emulated shared memory
    (compute-sanitizer is closed on the GPU pool, and GPU-side schedule fuzzing cannot see a missing
    __syncwarp).  The unmodified schedule must be race-free over several steps (one with a >= 1024) and reproduce
    the oracle's accumulator; with any ONE of its ten synchronisation points removed the detector must fire."""
    orc.build()
    exe = str(tmp_path / "emul_sched")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", "-o", exe, os.path.join(ROOT, "tests", "emul", "emul_sched.cpp"),
                           "-L" + os.path.join(ROOT, "oracle"), "-lorc", "-Wl,-rpath," + os.path.join(ROOT, "oracle")])
    out = subprocess.run([exe, "5", "0"], capture_output=True, text=True, timeout=300).stdout
    assert "races 0, accumulator words differing from the oracle 0" in out, out
    for site in range(1, 11):
        out = subprocess.run([exe, "2", str(site)], capture_output=True, text=True, timeout=300).stdout
        races = int(out.split("races ")[1].split(",")[0])
        assert races > 0, (site, out)


def test_cluster_kernel_race_detector_on_the_real_kernel_source(tmp_path):
    """tests/emul/emul_cl4s.cpp runs the REAL k_blind_rotate_cl4s (4 CTAs x 128 threads, B2_HOST_TRACE build) on
    512 host threads: distributed shared memory (mapa, st.async with mbarrier complete_tx), the mbarrier waits,
    __syncthreads, team barriers and cluster.sync are modelled by tests/emul/trace_rt.h with vector clocks.  Five
    CMUX steps (the double-buffered receive buffers are re-used twice) must be race-free and give the oracle's
    TRLWE; without any ONE of the five in-loop synchronisation points -- the two mbarrier waits included, which
    are not run as mutants on the GPU because they may hang it -- the detector must fire."""
    orc.build()
    exe = str(tmp_path / "emul_cl4s")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", "-o", exe, os.path.join(ROOT, "tests", "emul", "emul_cl4s.cpp"),
                           "-L" + os.path.join(ROOT, "oracle"), "-lorc", "-Wl,-rpath," + os.path.join(ROOT, "oracle")])
    out = subprocess.run([exe, "0"], capture_output=True, text=True, timeout=600).stdout
    assert "races 0, TRLWE words differing from the oracle 0" in out, out
    for site in range(1, 6):
        out = subprocess.run([exe, str(site)], capture_output=True, text=True, timeout=600).stdout
        races = int(out.split("races ")[1].split(",")[0])
        assert races > 0, (site, out)


def test_transform_kernels_race_detector_on_the_real_kernel_source(tmp_path):
    """tests/emul/emul_nttk.cpp runs the REAL k_ntt_fwd_s<0, true> / k_ntt_inv_s<true> (bulk-copy input prefetch:
    a staging buffer that is re-armed while it may still be read, an mbarrier with a transaction count) and
    k_polymul_s<2, 5, true> / <3, 3, true> (operands prefetched the same way) / k_cmux_s<4> as one CTA on host threads under
    the vector-clock tracer: no race in 2.5 M
    traced accesses, INTT(NTT(x)) == x, canonical residues, products equal to the schoolbook truth.  Without the
    barrier that keeps the producer from refilling a staging buffer (sites 11 / 12 / 13) the detector reports races
    and the deadlock that follows."""
    orc.build()
    exe = str(tmp_path / "emul_nttk")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", "-o", exe, os.path.join(ROOT, "tests", "emul", "emul_nttk.cpp"),
                           "-L" + os.path.join(ROOT, "oracle"), "-lorc", "-Wl,-rpath," + os.path.join(ROOT, "oracle")])
    out = subprocess.run([exe, "13", "0"], capture_output=True, text=True, timeout=900).stdout
    assert "races 0, round-trip words wrong 0, residues out of range 0, products / CMUX results wrong 0" in out, out
    for site in (11, 12, 13):
        out = subprocess.run([exe, "13", str(site), "5"], capture_output=True, text=True, timeout=900).stdout
        # the producer runs ahead: the phase completes twice before a consumer has waited once -> races on the buffer
        # and / or a consumer that waits for a parity that never comes again
        assert "DEADLOCK" in out or "races 0," not in out.splitlines()[-1], (site, out)


def test_whole_throughput_kernel_race_detector(tmp_path):
    """tests/emul/emul_brs.cpp: the REAL k_blind_rotate_s<4> as one CTA of 512 host threads -- four gate slots side by
    side, three of them occupied -- from table staging and pre-add / modulus switch / test vector through three
    CMUX steps to SampleExtract(0): no race (the gates' named barriers do not interfere, head and tail are ordered
    against the loop), lvl1 outputs equal to the oracle's; a dropped hand-over barrier is reported."""
    orc.build()
    exe = str(tmp_path / "emul_brs")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", "-o", exe, os.path.join(ROOT, "tests", "emul", "emul_brs.cpp"),
                           "-L" + os.path.join(ROOT, "oracle"), "-lorc", "-Wl,-rpath," + os.path.join(ROOT, "oracle")])
    out = subprocess.run([exe, "0"], capture_output=True, text=True, timeout=900).stdout
    assert "races 0, output words differing from the oracle 0" in out, out
    out = subprocess.run([exe, "9"], capture_output=True, text=True, timeout=900).stdout
    assert int(out.split("races ")[1].split(",")[0]) > 0, out


# -------------------------------------------------------------- host logic ---
def test_partition_covers_batch():
    from cufhe_b200.api import partition
    for count in (0, 1, 7, 16384, 2**18 + 3):
        for world in (1, 2, 3, 8):
            spans = [partition(count, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == count
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1


def test_abi_exports_every_declared_symbol():
    """The shared library loads and exports exactly what include/b2fhe.h declares."""
    from cufhe_b200 import _lib
    _lib.build()
    hdr = open(os.path.join(ROOT, "include", "b2fhe.h")).read()
    declared = set(re.findall(r"\b(b2fhe_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.SYMBOLS)
    L = C.CDLL(_lib.LIB_PATH)
    for s in declared:
        assert hasattr(L, s), s


def test_wire_formats_against_tfhepp_golden(keys):
    """SURVEY.md 8f rank 4.  tests/golden/wire_golden.json holds framing bytes and whole small
    archives written by the REAL TFHEpp's cereal serialisation (make_wire_golden.py); the
    engine's writer must reproduce them and its reader must accept them."""
    import json
    from cufhe_b200 import _lib
    _lib.build()
    L = _lib.lib()
    with open(os.path.join(ROOT, "tests", "golden", "wire_golden.json")) as f:
        G = json.load(f)
    BK, KSK = keys.bk.nbytes, keys.ksk.nbytes
    assert L.b2fhe_wire_evalkey_size() == G["evalkey_bytes"]
    blob = np.zeros(G["evalkey_bytes"], np.uint8)
    wrote = C.c_size_t()
    _lib.check(L.b2fhe_wire_evalkey_write(blob.ctypes.data, blob.size, keys.bk.ctypes.data, keys.ksk.ctypes.data, C.byref(wrote)))
    assert wrote.value == blob.size
    head = blob[:117].copy(); head[33:41] = 0
    assert head.tobytes().hex() == G["evalkey_head_hex"]
    assert blob[117 + BK:117 + BK + 48].tobytes().hex() == G["evalkey_mid_hex"]
    assert blob[117 + BK + 48 + KSK:].tobytes().hex() == G["evalkey_tail_hex"]
    ob, ok = C.c_size_t(), C.c_size_t()
    _lib.check(L.b2fhe_wire_evalkey_parse(blob.ctypes.data, blob.size, C.byref(ob), C.byref(ok)))
    assert (ob.value, ok.value) == (117, 117 + BK + 48)
    assert np.array_equal(blob[ob.value:ob.value + BK], keys.bk.view(np.uint8))
    assert np.array_equal(blob[ok.value:ok.value + KSK], keys.ksk.view(np.uint8))
    # truncation, trailing bytes, a key type the reader cannot skip
    assert L.b2fhe_wire_evalkey_parse(blob.ctypes.data, blob.size - 1, C.byref(ob), C.byref(ok)) == _lib.E_BAD_FORMAT
    more = np.concatenate([blob, np.zeros(1, np.uint8)])
    assert L.b2fhe_wire_evalkey_parse(more.ctypes.data, more.size, C.byref(ob), C.byref(ok)) == _lib.E_BAD_FORMAT
    odd = blob.copy(); odd[117 + BK:117 + BK + 4] = np.frombuffer((0x80000002).to_bytes(4, "little"), np.uint8)
    assert L.b2fhe_wire_evalkey_parse(odd.ctypes.data, odd.size, C.byref(ob), C.byref(ok)) == _lib.E_BAD_FORMAT
    assert b"cannot skip" in L.b2fhe_last_error_string()
    for level, words, dt in ((0, 637, np.uint16), (1, 1025, np.uint32)):
        ref = np.frombuffer(bytes.fromhex(G[f"tlwes_lvl{level}_hex"]), np.uint8)
        n, off = C.c_size_t(), C.c_size_t()
        _lib.check(L.b2fhe_wire_tlwes_parse(ref.ctypes.data, ref.size, level, C.byref(n), C.byref(off)))
        assert (n.value, off.value) == (2, 9)
        cts = ref[9:].view(dt).reshape(2, words).copy()
        mine = np.zeros(ref.size, np.uint8)
        _lib.check(L.b2fhe_wire_tlwes_write(mine.ctypes.data, mine.size, level, cts.ctypes.data, 2, None))
        assert np.array_equal(mine, ref)


def test_product_path_never_touches_oracle():
    """Nothing under cufhe_b200/ or include/ imports, links or calls the oracle."""
    for top in ("cufhe_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".inl", ".h", ".hpp")):
                    src = open(os.path.join(dirpath, f)).read()
                    for needle in ("import orc", "from oracle", "liborc", "orc_", "cufhe_oracle"):
                        assert needle not in src, (f, needle)
