"""Build + ctypes binding of the C-ABI shared library (include/b2fhe.h).

The library is the product: there is no Python or CPU fallback.  If it cannot be
loaded, every call raises.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "libb2fhe.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "--shared", "-Xcompiler", "-fPIC",
]

# every symbol include/b2fhe.h declares
SYMBOLS = [
    "b2fhe_init", "b2fhe_setup_devices", "b2fhe_init_device", "b2fhe_adopt_keys", "b2fhe_key_pointers", "b2fhe_cleanup",
    "b2fhe_num_devices", "b2fhe_last_error_string", "b2fhe_gate_batch", "b2fhe_gate_batch_mixed",
    "b2fhe_gate_batch_host", "b2fhe_wire_evalkey_size", "b2fhe_wire_evalkey_write", "b2fhe_wire_evalkey_parse",
    "b2fhe_init_from_evalkey", "b2fhe_wire_tlwes_size", "b2fhe_wire_tlwes_write", "b2fhe_wire_tlwes_parse",
    "b2fhe_key_cache_size", "b2fhe_key_cache_export", "b2fhe_init_from_key_cache", "b2fhe_synchronize", "b2fhe_gate_batch_ptrs", "b2fhe_defer_gate", "b2fhe_defer_h2d", "b2fhe_defer_d2h",
    "b2fhe_defer_flush", "b2fhe_defer_fence", "b2fhe_defer_fence_buffer", "b2fhe_defer_query", "b2fhe_defer_query_stream", "b2fhe_defer_pending", "b2fhe_blind_rotate_batch", "b2fhe_bootstrap_batch", "b2fhe_keyswitch_batch",
    "b2fhe_bootstrap_tlwe2trlwe_batch", "b2fhe_sample_extract_keyswitch_batch", "b2fhe_refresh_batch", "b2fhe_trgsw2ntt_batch", "b2fhe_cmux_ntt_batch",
    "b2fhe_ntt1024_batch", "b2fhe_intt1024_batch", "b2fhe_polymul1024_batch", "b2fhe_polymul1024_exact_batch", "b2fhe_set_latency_mode",
    "b2fhe_gate_batch_noalloc", "b2fhe_noalloc_max_gates",
    "b2fhe_set_gates_per_cta", "b2fhe_set_br_variant", "b2fhe_set_ks_variant", "b2fhe_kernel_launches", "b2fhe_int_peak",
]


def _sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(ROOT, "include", "b2fhe.h")]


def build(force=False, verbose=False):
    """Compile csrc/ into libb2fhe.so for sm_100a (nvcc cross-compiles without a GPU)."""
    if not force and os.path.exists(LIB_PATH):
        newest = max(os.path.getmtime(s) for s in _sources())
        if os.path.getmtime(LIB_PATH) >= newest:
            return LIB_PATH
    cmd = [NVCC] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [
        "-o", LIB_PATH, os.path.join(CSRC, "b2fhe_api.cu")]
    subprocess.check_call(cmd)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i32, sz = C.c_void_p, C.c_int, C.c_size_t
    L.b2fhe_init.argtypes = [i32, vp, vp]
    L.b2fhe_setup_devices.argtypes = [i32]
    L.b2fhe_init_device.argtypes = [i32, vp, vp]
    L.b2fhe_adopt_keys.argtypes = [i32, vp, vp]
    L.b2fhe_key_pointers.argtypes = [i32, C.POINTER(vp), C.POINTER(vp)]
    L.b2fhe_last_error_string.restype = C.c_char_p
    L.b2fhe_gate_batch.argtypes = [i32, vp, i32, i32, vp, vp, vp, vp, sz]
    L.b2fhe_gate_batch_mixed.argtypes = [i32, vp, vp, i32, vp, vp, vp, vp, sz]
    L.b2fhe_gate_batch_host.argtypes = [i32, vp, i32, vp, i32, vp, vp, vp, vp, sz]
    L.b2fhe_gate_batch_ptrs.argtypes = [i32, vp, vp, i32, vp, vp, vp, vp, sz]
    L.b2fhe_wire_evalkey_size.restype = sz
    L.b2fhe_wire_evalkey_size.argtypes = []
    L.b2fhe_wire_evalkey_write.argtypes = [vp, sz, vp, vp, C.POINTER(sz)]
    L.b2fhe_wire_evalkey_parse.argtypes = [vp, sz, C.POINTER(sz), C.POINTER(sz)]
    L.b2fhe_init_from_evalkey.argtypes = [i32, vp, sz]
    L.b2fhe_wire_tlwes_size.restype = sz
    L.b2fhe_wire_tlwes_size.argtypes = [i32, sz]
    L.b2fhe_wire_tlwes_write.argtypes = [vp, sz, i32, vp, sz, C.POINTER(sz)]
    L.b2fhe_wire_tlwes_parse.argtypes = [vp, sz, i32, C.POINTER(sz), C.POINTER(sz)]
    L.b2fhe_key_cache_size.restype = sz
    L.b2fhe_key_cache_size.argtypes = []
    L.b2fhe_key_cache_export.argtypes = [i32, vp, sz, C.POINTER(sz)]
    L.b2fhe_init_from_key_cache.argtypes = [i32, vp, sz]
    L.b2fhe_defer_gate.argtypes = [i32, vp, i32, i32, vp, vp, vp, vp]
    L.b2fhe_defer_d2h.argtypes = [i32, vp, vp, sz]
    L.b2fhe_defer_h2d.argtypes = [i32, vp, vp, sz]
    L.b2fhe_defer_flush.argtypes = [i32]
    L.b2fhe_defer_fence.argtypes = [i32, vp]
    L.b2fhe_defer_fence_buffer.argtypes = [i32, vp, vp]
    L.b2fhe_defer_query.argtypes = [i32, C.POINTER(i32)]
    L.b2fhe_defer_query_stream.argtypes = [i32, vp, C.POINTER(i32)]
    L.b2fhe_defer_pending.argtypes = [i32]
    L.b2fhe_blind_rotate_batch.argtypes = [i32, vp, vp, vp, C.c_uint32, sz]
    L.b2fhe_bootstrap_batch.argtypes = [i32, vp, vp, vp, C.c_uint32, sz]
    L.b2fhe_keyswitch_batch.argtypes = [i32, vp, vp, vp, sz]
    L.b2fhe_bootstrap_tlwe2trlwe_batch.argtypes = [i32, vp, vp, vp, C.c_uint32, sz]
    L.b2fhe_sample_extract_keyswitch_batch.argtypes = [i32, vp, vp, vp, sz]
    L.b2fhe_refresh_batch.argtypes = [i32, vp, vp, vp, C.c_uint32, sz]
    L.b2fhe_trgsw2ntt_batch.argtypes = [i32, vp, vp, vp, sz]
    L.b2fhe_cmux_ntt_batch.argtypes = [i32, vp, vp, vp, vp, vp, sz, i32]
    L.b2fhe_ntt1024_batch.argtypes = [i32, vp, vp, vp, sz]
    L.b2fhe_intt1024_batch.argtypes = [i32, vp, vp, vp, sz]
    L.b2fhe_polymul1024_batch.argtypes = [i32, vp, vp, vp, vp, sz]
    L.b2fhe_polymul1024_exact_batch.argtypes = [i32, vp, vp, vp, vp, sz]
    L.b2fhe_set_latency_mode.argtypes = [i32]
    L.b2fhe_gate_batch_noalloc.argtypes = [i32, vp, i32, vp, vp, vp, sz]
    L.b2fhe_noalloc_max_gates.restype = sz
    L.b2fhe_noalloc_max_gates.argtypes = []
    L.b2fhe_set_gates_per_cta.argtypes = [i32]
    L.b2fhe_set_br_variant.argtypes = [i32]
    L.b2fhe_set_ks_variant.argtypes = [i32]
    L.b2fhe_kernel_launches.restype = C.c_uint64
    L.b2fhe_int_peak.argtypes = [i32, i32, C.POINTER(C.c_double)]
    _lib = L
    return L


class B2FheError(RuntimeError):
    pass


E_NOT_INITIALIZED, E_BAD_ARGUMENT, E_NO_DEVICE, E_BAD_FORMAT, E_TOO_LARGE = 10001, 10002, 10003, 10004, 10005   # include/b2fhe.h


def check(code):
    if code != 0:
        msg = lib().b2fhe_last_error_string()
        raise B2FheError(f"b2fhe error {code}: {msg.decode() if msg else ''}")
