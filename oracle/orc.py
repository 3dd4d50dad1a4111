"""ctypes binding of the CPU oracle (oracle/cufhe_oracle.c).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(cufhe_b200/) must never import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liborc.so")

n, N = 636, 1024
LVL0_WORDS, LVL1_WORDS = n + 1, N + 1
BK_WORDS = n * 4 * 2 * N
KSK_WORDS = N * 7 * 2 * LVL0_WORDS
MU0, MU1 = 1 << 13, 1 << 29
OPS = ["NAND", "OR", "ORYN", "ORNY", "AND", "ANDYN", "ANDNY", "NOR", "XOR", "XNOR",
       "MUX", "NMUX", "NOT", "COPY"]
OP = {name: i for i, name in enumerate(OPS)}


def build(force=False):
    src = os.path.join(_HERE, "cufhe_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liborc.so"], stdout=subprocess.DEVNULL)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        _lib = C.CDLL(_LIB)
        _lib.orc_decrypt_lvl0.restype = C.c_int
        _lib.orc_decrypt_lvl1.restype = C.c_int
        _lib.orc_truth.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Keys:
    """Seeded key set in TFHEpp's raw array layouts (a4, a6)."""

    def __init__(self, seed=20261104):
        self.seed = seed
        self.sk0 = np.zeros(n, np.int32)
        self.sk1 = np.zeros(N, np.int32)
        self.bk = np.zeros(BK_WORDS, np.uint32)
        self.ksk = np.zeros(KSK_WORDS, np.uint16)
        lib().orc_keygen(C.c_uint64(seed), _p(self.sk0), _p(self.sk1), _p(self.bk), _p(self.ksk))


def encrypt(keys, bits, level, base_index=0, seed=None):
    bits = np.asarray(bits, np.uint8).ravel()
    seed = keys.seed + 1 if seed is None else seed
    if level == 0:
        out = np.zeros((len(bits), LVL0_WORDS), np.uint16)
        for i, b in enumerate(bits):
            lib().orc_encrypt_lvl0(C.c_uint64(seed), C.c_uint64(base_index + i), int(b), _p(keys.sk0), _p(out[i]))
    else:
        out = np.zeros((len(bits), LVL1_WORDS), np.uint32)
        for i, b in enumerate(bits):
            lib().orc_encrypt_lvl1(C.c_uint64(seed), C.c_uint64(base_index + i), int(b), _p(keys.sk1), _p(out[i]))
    return out


def decrypt(keys, cts, level):
    cts = np.ascontiguousarray(cts)
    if level == 0:
        # vectorised tlweSymDecrypt (TFHEpp tlwe.hpp:77-103)
        ph = cts[:, n].astype(np.int64) - (cts[:, :n].astype(np.int64) @ keys.sk0.astype(np.int64))
        ph = (ph & 0xFFFF).astype(np.uint16).astype(np.int16)
    else:
        ph = cts[:, N].astype(np.int64) - (cts[:, :N].astype(np.int64) @ keys.sk1.astype(np.int64))
        ph = (ph & 0xFFFFFFFF).astype(np.uint32).astype(np.int32)
    return (ph > 0).astype(np.uint8)


def truth(op, a, b=0, c=0):
    return lib().orc_truth(int(op), int(a), int(b), int(c))


def gate_batch(keys, op, level, in0, in1=None, in2=None, ops=None, threads=0):
    in0 = np.ascontiguousarray(in0)
    out = np.zeros_like(in0)
    in1 = None if in1 is None else np.ascontiguousarray(in1)
    in2 = None if in2 is None else np.ascontiguousarray(in2)
    ops_a = None if ops is None else np.ascontiguousarray(ops, dtype=np.uint8)
    lib().orc_gate_batch(int(op), _p(ops_a), int(level), _p(out), _p(in0), _p(in1), _p(in2),
                         C.c_size_t(in0.shape[0]), _p(keys.bk), _p(keys.ksk), int(threads))
    return out


def blind_rotate_extract(keys, c, mu=MU1, schoolbook=False):
    c = np.ascontiguousarray(c, dtype=np.uint16)
    out = np.zeros(LVL1_WORDS, np.uint32)
    lib().orc_blind_rotate_extract(_p(out), _p(c), C.c_uint32(mu), _p(keys.bk), int(schoolbook))
    return out


def blind_rotate_trlwe(keys, c, mu=MU1):
    c = np.ascontiguousarray(c, dtype=np.uint16)
    out = np.zeros(2 * N, np.uint32)
    lib().orc_blind_rotate_trlwe(_p(out), _p(c), C.c_uint32(mu), _p(keys.bk))
    return out


def sample_extract_keyswitch(keys, trlwe):
    t = np.ascontiguousarray(trlwe, dtype=np.uint32)
    out = np.zeros(LVL0_WORDS, np.uint16)
    lib().orc_sample_extract_keyswitch(_p(out), _p(t), _p(keys.ksk))
    return out


def refresh(keys, trlwe, mu=MU1):
    t = np.ascontiguousarray(trlwe, dtype=np.uint32)
    out = np.zeros(2 * N, np.uint32)
    lib().orc_refresh(_p(out), _p(t), C.c_uint32(mu), _p(keys.bk), _p(keys.ksk))
    return out


def keyswitch(keys, x):
    x = np.ascontiguousarray(x, dtype=np.uint32)
    out = np.zeros(LVL0_WORDS, np.uint16)
    lib().orc_keyswitch(_p(out), _p(x), _p(keys.ksk))
    return out


def cmux_step(acc, abar, bk_i, schoolbook=False):
    acc = np.array(acc, dtype=np.uint32, copy=True).reshape(2 * N)
    bk_i = np.ascontiguousarray(bk_i, dtype=np.uint32)
    lib().orc_cmux_step(_p(acc), C.c_uint32(abar), _p(bk_i), int(schoolbook))
    return acc.reshape(2, N)


def cmux(trgsw, c1, c0, schoolbook=False):
    """CMUXNTT: c0 + (c1 - c0) [.] trgsw on TRLWEs (2 x 1024 u32); trgsw raw [4][2][1024]."""
    trgsw = np.ascontiguousarray(trgsw, dtype=np.uint32)
    c1 = np.ascontiguousarray(c1, dtype=np.uint32)
    c0 = np.ascontiguousarray(c0, dtype=np.uint32)
    out = np.zeros(2 * N, np.uint32)
    lib().orc_cmux(_p(out), _p(trgsw), _p(c1), _p(c0), int(schoolbook))
    return out


def polymul(a, b, schoolbook=True):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    b = np.ascontiguousarray(b, dtype=np.uint32)
    out = np.zeros(N, np.uint32)
    (lib().orc_polymul_schoolbook if schoolbook else lib().orc_polymul_ntt)(_p(out), _p(a), _p(b))
    return out
