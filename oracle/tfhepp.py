"""ctypes binding of oracle/_ref/libtfhepp_ref.so -- the REFERENCE's CPU library (TFHEpp)
compiled from /root/reference by oracle/Makefile.  TEST INFRASTRUCTURE ONLY (see
oracle/tfhepp_ref.cpp).  available() is False when the build output is absent."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_ref", "libtfhepp_ref.so")
n, N = 636, 1024
_lib = None


def available():
    return os.path.exists(LIB)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB)
        L.ref_bk_bytes.restype = C.c_size_t
        L.ref_ksk_bytes.restype = C.c_size_t
        L.ref_bench_nand.restype = C.c_double
        L.ref_bench_nand.argtypes = [C.c_int, C.c_int]
        _lib = L
    return _lib


class Keys:
    """A TFHEpp-generated key set exported in TFHEpp's raw layouts (same attribute names
    as oracle.orc.Keys, so either can drive the oracle or the CUDA engine)."""

    def __init__(self):
        L = lib()
        L.ref_keygen()
        assert L.ref_bk_bytes() == n * 4 * 2 * N * 4 and L.ref_ksk_bytes() == N * 7 * 2 * (n + 1) * 2
        self.sk0 = np.zeros(n, np.int32)
        self.sk1 = np.zeros(N, np.int32)
        self.bk = np.zeros(n * 4 * 2 * N, np.uint32)
        self.ksk = np.zeros(N * 7 * 2 * (n + 1), np.uint16)
        L.ref_export_keys(self.sk0.ctypes.data_as(C.c_void_p), self.sk1.ctypes.data_as(C.c_void_p),
                          self.bk.ctypes.data_as(C.c_void_p), self.ksk.ctypes.data_as(C.c_void_p))
        self.seed = 0


def encrypt(bits, level):
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    out = np.zeros((len(bits), n + 1), np.uint16) if level == 0 else np.zeros((len(bits), N + 1), np.uint32)
    lib().ref_encrypt(int(level), bits.ctypes.data_as(C.c_void_p), len(bits), out.ctypes.data_as(C.c_void_p))
    return out


def decrypt(cts, level):
    cts = np.ascontiguousarray(cts)
    bits = np.zeros(cts.shape[0], np.uint8)
    lib().ref_decrypt(int(level), cts.ctypes.data_as(C.c_void_p), cts.shape[0], bits.ctypes.data_as(C.c_void_p))
    return bits


def homnand_lvl0(x0, x1, threads=0):
    x0, x1 = np.ascontiguousarray(x0), np.ascontiguousarray(x1)
    out = np.zeros_like(x0)
    lib().ref_homnand_lvl0(out.ctypes.data_as(C.c_void_p), x0.ctypes.data_as(C.c_void_p),
                           x1.ctypes.data_as(C.c_void_p), x0.shape[0], threads or os.cpu_count())
    return out


# ---- wire-format pins (TFHEpp's own cereal portable_binary serialisation) -------------------
def _wire_lib():
    L = lib()
    L.ref_serialize_evalkey.restype = C.c_size_t
    L.ref_serialize_evalkey.argtypes = [C.c_int, C.c_void_p, C.c_size_t]
    L.ref_deserialize_evalkey.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    L.ref_serialize_tlwes.restype = C.c_size_t
    L.ref_serialize_tlwes.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
    L.ref_deserialize_tlwes.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
    return L


def serialize_evalkey(which=0):
    """EvalKey archive written by TFHEpp (which=1 also holds the FFT form of the key)."""
    L = _wire_lib()
    nbytes = L.ref_serialize_evalkey(which, None, 0)
    buf = np.zeros(nbytes, np.uint8)
    L.ref_serialize_evalkey(which, buf.ctypes.data, nbytes)
    return buf


def deserialize_evalkey(blob):
    """Loads an EvalKey archive with TFHEpp's own reader -> (status, bk, ksk) raw arrays."""
    L = _wire_lib()
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    bk = np.zeros(n * 4 * 2 * N, np.uint32)
    ksk = np.zeros(N * 7 * 2 * (n + 1), np.uint16)
    rc = L.ref_deserialize_evalkey(blob.ctypes.data, blob.size, bk.ctypes.data, ksk.ctypes.data)
    return rc, bk, ksk


def serialize_tlwes(cts, level):
    L = _wire_lib()
    cts = np.ascontiguousarray(cts)
    nbytes = L.ref_serialize_tlwes(level, cts.ctypes.data, cts.shape[0], None, 0)
    buf = np.zeros(nbytes, np.uint8)
    L.ref_serialize_tlwes(level, cts.ctypes.data, cts.shape[0], buf.ctypes.data, nbytes)
    return buf


def deserialize_tlwes(blob, level, cap):
    L = _wire_lib()
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    out = np.zeros((cap, n + 1), np.uint16) if level == 0 else np.zeros((cap, N + 1), np.uint32)
    cnt = L.ref_deserialize_tlwes(level, blob.ctypes.data, blob.size, out.ctypes.data, cap)
    return cnt, out
