"""Byte parity with the REFERENCE cuFHE GPU library itself.

oracle/_ref/libcufhe_ref.so is the reference's own CUDA sources (src/*.cu, unmodified,
compiled for sm_100 by oracle/Makefile) behind a thin C wrapper (oracle/cufhe_ref_gpu.cu).
On the same keys and the same input ciphertexts, for all 14 ops and both I/O levels:

    reference cuFHE GPU output  ==  CPU oracle output  ==  our engine's output   (bytes)

This is the pin the oracle's header promises (the reference ships no golden
vectors; its device code is the specification)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from oracle import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libcufhe_ref.so")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(LIB), reason="oracle/_ref/libcufhe_ref.so not built")]


@pytest.fixture(scope="module")
def ref(keys):
    L = C.CDLL(LIB)
    L.cufheref_bench_nand.restype = C.c_double
    L.cufheref_bench_nand.argtypes = [C.c_int, C.c_int]
    assert L.cufheref_init(keys.bk.ctypes.data_as(C.c_void_p), keys.ksk.ctypes.data_as(C.c_void_p)) == 0
    yield L
    L.cufheref_cleanup()


def ref_gate(L, op, level, x0, x1, x2):
    out = np.zeros_like(x0)
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)  # noqa: E731
    assert L.cufheref_gate(int(op), int(level), p(out), p(x0), p(x1), p(x2), x0.shape[0]) == 0
    return out


@pytest.mark.parametrize("level", [0, 1])
def test_reference_gpu_equals_oracle_equals_engine(engine, keys, ref, level):
    cf = engine
    cnt, nchk = 64, 6          # 64 gates per op against the reference GPU build; the (slow) oracle on the first 6
    rng = np.random.default_rng(40 + level)
    st = cf.Stream(0).Create()
    for name in orc.OPS:
        op = orc.OP[name]
        b = [rng.integers(0, 2, cnt) for _ in range(3)]
        x = [orc.encrypt(keys, b[k], level, base_index=70000 + 1000 * op + 100 * k) for k in range(3)]
        # ragged / edge inputs go to the reference too: all-zero ciphertext, maximal mask, all-ones
        x[0][3, :] = 0
        x[1][4, :-1] = 0xFFFF if level == 0 else 0xFFFFFFFF
        x[0][5, :] = 0xFFFF if level == 0 else 0xFFFFFFFF
        unary, ternary = name in ("NOT", "COPY"), name in ("MUX", "NMUX")
        x1 = None if unary else x[1]
        x2 = x[2] if ternary else None
        got_ref = ref_gate(ref, op, level, x[0], x1, x2)
        want = orc.gate_batch(keys, op, level, x[0][:nchk], None if x1 is None else x1[:nchk], None if x2 is None else x2[:nchk])
        assert np.array_equal(got_ref[:nchk], want), f"{name} lvl{level}: oracle differs from the reference cuFHE GPU build"
        cts = [cf.Ctxt(cnt, level).set(v) for v in x]
        out = cf.Ctxt(cnt, level)
        for c in cts:
            cf.CtxtCopyH2D(c, st)
        cf.api._gate_device(name, out, cts[0], None if unary else cts[1], cts[2] if ternary else None, st)
        cf.CtxtCopyD2H(out, st)
        cf.Synchronize()
        assert np.array_equal(out.tlwehost, got_ref), f"{name} lvl{level}: engine differs from the reference cuFHE GPU build"


def test_reference_gpu_throughput_recorded(ref):
    """Times the reference build with its own method (streams = 4 x SMs) and leaves the number in
    gpurun_out/ for BASELINE.md; no assertion on speed."""
    import torch
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    count = 8 * sms
    secs = ref.cufheref_bench_nand(count, 4 * sms)
    lat = min(ref.cufheref_bench_nand(1, 1) for _ in range(5))      # one gate, one stream, copies included
    rec = {"reference_cufhe_gpu_nand_lvl0": {"gates": count, "streams": 4 * sms, "seconds": secs,
                                             "gates_per_s": count / secs, "ms_per_gate_throughput": secs / count * 1e3,
                                             "single_gate_latency_ms": lat * 1e3}}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "reference_gpu_bench.json"), "w") as f:
        json.dump(rec, f)
    print(json.dumps(rec))
    assert secs > 0
