"""Levelised gate circuits over device-resident ciphertext batches (SURVEY.md 8f rank 1,
BASELINE config 5).

The reference has no circuit harness; its consumers chain the device-resident `g*` gates
(include/cufhe_gpu.cuh:229-252) one ciphertext at a time.  Here a circuit is evaluated
for many independent instances at once: every level is ONE batched launch over all
instances (and over all bit positions that are ready), ciphertexts never leave the GPU,
and instances are partitioned across GPUs so no level needs inter-GPU traffic.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .api import OP, _torch_dtype, _words


class Wires:
    """`n` wires x `instances` ciphertexts on one device: tensor [n][instances][words].
    A wire (or a run of consecutive wires) is a contiguous batch for the C ABI."""

    def __init__(self, n, instances, level=0, device=0):
        self.n, self.instances, self.level, self.device = int(n), int(instances), int(level), int(device)
        self.t = torch.empty((self.n, self.instances, _words(level)), dtype=_torch_dtype(level),
                             device=torch.device("cuda", self.device))

    def ptr(self, wire=0):
        return self.t.data_ptr() + wire * self.instances * self.t.shape[2] * self.t.element_size()

    def load(self, host_array):
        """host_array: [n][instances][words] numpy (TFHEpp raw ciphertexts)."""
        a = np.ascontiguousarray(host_array).reshape(self.t.shape)
        self.t.copy_(torch.from_numpy(a.view(np.int16 if self.level == 0 else np.int32)))
        return self

    def numpy(self):
        return self.t.cpu().numpy().view(np.uint16 if self.level == 0 else np.uint32)


def _gate(op, out, out_wire, a, a_wire, b, b_wire, st, wires=1):
    """out[out_wire : out_wire+wires] = op(a[a_wire : ...], b[b_wire : ...]) -- one launch."""
    _lib.check(_lib.lib().b2fhe_gate_batch(
        out.device, st.st(), OP[op], out.level, C.c_void_p(out.ptr(out_wire)), C.c_void_p(a.ptr(a_wire)),
        C.c_void_p(b.ptr(b_wire)), None, wires * out.instances))
    return wires * out.instances


def ripple_carry_add(a, b, st, scratch=None):
    """sum = a + b mod 2^n for `instances` independent n-bit additions (wire i = bit i, LSB
    first).  Levels:
      1. x_i = a_i XOR b_i, g_i = a_i AND b_i for ALL i          (2 launches of n*instances)
      2. carry chain, i = 1..n-2: t = c_i AND x_i ; c_{i+1} = g_i OR t   (2 launches per bit)
      3. s_i = x_i XOR c_i for all i >= 1                         (1 launch)
    Returns (sum Wires, number of gate bootstraps evaluated)."""
    n, inst = a.n, a.instances
    assert (b.n, b.instances, b.level, b.device) == (n, inst, a.level, a.device)
    x = Wires(n, inst, a.level, a.device)
    g = Wires(n, inst, a.level, a.device)
    c = Wires(n, inst, a.level, a.device)       # c[i] = carry INTO bit i (c[0] unused)
    t = Wires(1, inst, a.level, a.device)
    s = Wires(n, inst, a.level, a.device)
    # the temporaries are allocated on torch's CURRENT stream but used by kernels on `st`: tell the caching
    # allocator, or their blocks could be reused while those kernels still run
    if st._st is not None and st._st != torch.cuda.current_stream(st._st.device):
        for w in (x, g, c, t, s):
            w.t.record_stream(st._st)
    gates = 0
    gates += _gate("XOR", x, 0, a, 0, b, 0, st, wires=n)
    gates += _gate("AND", g, 0, a, 0, b, 0, st, wires=n - 1)     # g_{n-1} is never needed
    # carry into bit 1 is g_0
    _lib.check(_lib.lib().b2fhe_gate_batch(a.device, st.st(), OP["COPY"], a.level, C.c_void_p(c.ptr(1)),
                                           C.c_void_p(g.ptr(0)), None, None, inst))
    for i in range(1, n - 1):
        gates += _gate("AND", t, 0, c, i, x, i, st)
        gates += _gate("OR", c, i + 1, g, i, t, 0, st)
    _lib.check(_lib.lib().b2fhe_gate_batch(a.device, st.st(), OP["COPY"], a.level, C.c_void_p(s.ptr(0)),
                                           C.c_void_p(x.ptr(0)), None, None, inst))
    gates += _gate("XOR", s, 1, x, 1, c, 1, st, wires=n - 1)
    return s, gates


def adder_gate_count(n):
    """bootstraps per instance of ripple_carry_add."""
    return n + (n - 1) + 2 * (n - 2) + (n - 1)
