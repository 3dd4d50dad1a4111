"""Host-side mirror of cuFHE's public API for the gate-bootstrapping path.

Same names and argument meaning as the reference's `namespace cufhe`
(include/cufhe_gpu.cuh:42-315, src/cufhe_gates_gpu.cu): SetGPUNum, Initialize,
CleanUp, Synchronize, Stream, StreamQuery, Ctxt, And ... NMux, Not, Copy and the
device-resident g* twins -- but a `Ctxt` here is a BATCH of ciphertexts, so one call
is one batched launch.  Everything below the ctypes call is CUDA (libb2fhe.so);
torch is used only for device memory, streams and pinned host buffers.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

OPS = ["NAND", "OR", "ORYN", "ORNY", "AND", "ANDYN", "ANDNY", "NOR", "XOR", "XNOR",
       "MUX", "NMUX", "NOT", "COPY"]
OP = {n: i for i, n in enumerate(OPS)}
LVL0_WORDS, LVL1_WORDS = 637, 1025
BK_RAW_WORDS = 636 * 4 * 2 * 1024
KSK_RAW_WORDS = 1024 * 7 * 2 * 637
BK_NTT_BYTES = 636 * 4 * 2 * 2 * 1024 * 4
KSK_DEV_BYTES = 1024 * 7 * 3 * 320 * 4

_gpu_num = 1
_devices = []          # CUDA ordinals this process drives
_stream_count = 0


def _np_dtype(level):
    return np.uint16 if level == 0 else np.uint32


def _torch_dtype(level):
    return torch.int16 if level == 0 else torch.int32


def _words(level):
    return LVL0_WORDS if level == 0 else LVL1_WORDS


def SetGPUNum(gpu_num):
    """cufhe::SetGPUNum (src/cufhe_gates_gpu.cu:38)."""
    global _gpu_num
    _gpu_num = int(gpu_num)


def Initialize(bk_raw, ksk_raw, device=None):
    """cufhe::Initialize(EvalKey) (src/cufhe_gates_gpu.cu:42-47).

    bk_raw / ksk_raw: TFHEpp's raw BootstrappingKey<lvl01param> (uint32) and
    KeySwitchingKey<lvl10param> (uint16) arrays.  device=None: single process driving
    GPUs 0.._gpu_num-1 (the reference's model).  device=k: one process per GPU.
    """
    global _devices
    bk = np.ascontiguousarray(bk_raw, dtype=np.uint32).ravel()
    ksk = np.ascontiguousarray(ksk_raw, dtype=np.uint16).ravel()
    if bk.size != BK_RAW_WORDS or ksk.size != KSK_RAW_WORDS:
        raise ValueError("key arrays have the wrong size for the default 128-bit parameter set")
    L = _lib.lib()
    if device is None:
        _lib.check(L.b2fhe_init(_gpu_num, bk.ctypes.data, ksk.ctypes.data))
        _devices = list(range(_gpu_num))
    else:
        _lib.check(L.b2fhe_init_device(int(device), bk.ctypes.data, ksk.ctypes.data))
        _devices = [int(device)]


def InitializeFromArchive(blob):
    """Initialize from a serialised TFHEpp::EvalKey (cereal portable_binary, e.g. the TFHEpp
    tutorial's cloud.key; TFHEpp/include/cloudkey.hpp:103-107) holding bk<lvl01> + iksk<lvl10>."""
    global _devices
    blob = np.ascontiguousarray(np.frombuffer(blob, dtype=np.uint8) if isinstance(blob, (bytes, bytearray, memoryview)) else blob,
                                dtype=np.uint8)
    _lib.check(_lib.lib().b2fhe_init_from_evalkey(_gpu_num, blob.ctypes.data, blob.size))
    _devices = list(range(_gpu_num))


def SerializeEvalKey(bk_raw, ksk_raw):
    """The EvalKey archive TFHEpp itself would write for these raw keys (numpy uint8)."""
    bk = np.ascontiguousarray(bk_raw, dtype=np.uint32).ravel()
    ksk = np.ascontiguousarray(ksk_raw, dtype=np.uint16).ravel()
    if bk.size != BK_RAW_WORDS or ksk.size != KSK_RAW_WORDS:
        raise ValueError("key arrays have the wrong size for the default 128-bit parameter set")
    L = _lib.lib()
    blob = np.zeros(L.b2fhe_wire_evalkey_size(), np.uint8)
    _lib.check(L.b2fhe_wire_evalkey_write(blob.ctypes.data, blob.size, bk.ctypes.data, ksk.ctypes.data, None))
    return blob


def ExportKeyCache(device=0):
    """This engine's CONVERTED keys of `device` as one checksummed blob (numpy uint8)."""
    L = _lib.lib()
    blob = np.zeros(L.b2fhe_key_cache_size(), np.uint8)
    _lib.check(L.b2fhe_key_cache_export(int(device), blob.ctypes.data, blob.size, None))
    return blob


def InitializeFromKeyCache(blob):
    """Initialize without converting: uploads an ExportKeyCache blob and replicates it."""
    global _devices
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    _lib.check(_lib.lib().b2fhe_init_from_key_cache(_gpu_num, blob.ctypes.data, blob.size))
    _devices = list(range(_gpu_num))


def SerializeCtxts(cts, level):
    """std::vector<TLWE<P>> archive of a (count, words) array (TFHEpp tutorial cloud.data)."""
    cts = np.ascontiguousarray(cts, dtype=_np_dtype(level))
    L = _lib.lib()
    blob = np.zeros(L.b2fhe_wire_tlwes_size(level, cts.shape[0]), np.uint8)
    _lib.check(L.b2fhe_wire_tlwes_write(blob.ctypes.data, blob.size, level, cts.ctypes.data, cts.shape[0], None))
    return blob


def DeserializeCtxts(blob, level):
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    n, off = C.c_size_t(), C.c_size_t()
    _lib.check(_lib.lib().b2fhe_wire_tlwes_parse(blob.ctypes.data, blob.size, level, C.byref(n), C.byref(off)))
    words = LVL0_WORDS if level == 0 else LVL1_WORDS
    return blob[off.value:].view(_np_dtype(level)).reshape(n.value, words).copy()


def broadcast_key_buffers(bk_t, ksk_t, src=0, group=None):
    """The one collective of the path: the converted bootstrapping key and the key-switching
    key go from rank `src` to every rank (NCCL over NVLink for CUDA tensors; gloo in the CPU
    tests).  Steady state has no collective: gates are independent."""
    import torch.distributed as dist
    dist.broadcast(bk_t, src=src, group=group)
    dist.broadcast(ksk_t, src=src, group=group)
    return bk_t, ksk_t


def InitializeBroadcast(bk_raw, ksk_raw, device, group=None, src=0):
    """One process per GPU: rank `src` converts the keys, every other rank receives the
    NTT-domain key and the key-switching key by ONE broadcast each (NCCL over NVLink for
    CUDA tensors) and adopts them.  Replaces the reference's per-GPU host loop
    (src/bootstrap_gpu.cu:116-137, src/keyswitch_gpu.cu:10-15).  Returns the tensors
    that own the adopted key memory (keep them alive until CleanUp)."""
    import torch.distributed as dist
    global _devices
    L = _lib.lib()
    rank = dist.get_rank(group)
    dev = torch.device("cuda", int(device))
    if rank == src:
        Initialize(bk_raw, ksk_raw, device=device)
        pb, pk = C.c_void_p(), C.c_void_p()
        _lib.check(L.b2fhe_key_pointers(int(device), C.byref(pb), C.byref(pk)))
        bk_t = _wrap_device_bytes(pb.value, BK_NTT_BYTES, dev)
        ksk_t = _wrap_device_bytes(pk.value, KSK_DEV_BYTES, dev)
        broadcast_key_buffers(bk_t, ksk_t, src, group)
        return bk_t, ksk_t
    bk_t = torch.empty(BK_NTT_BYTES, dtype=torch.uint8, device=dev)
    ksk_t = torch.empty(KSK_DEV_BYTES, dtype=torch.uint8, device=dev)
    broadcast_key_buffers(bk_t, ksk_t, src, group)
    torch.cuda.synchronize(dev)
    _lib.check(L.b2fhe_adopt_keys(int(device), bk_t.data_ptr(), ksk_t.data_ptr()))
    _devices = [int(device)]
    return bk_t, ksk_t


def _wrap_device_bytes(ptr, nbytes, dev):
    """A torch uint8 view over library-owned device memory (no copy)."""
    class _Holder:
        pass
    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}
    return torch.as_tensor(h, device=dev)


def CleanUp():
    """cufhe::CleanUp (src/cufhe_gates_gpu.cu:49-53)."""
    global _devices
    _lib.check(_lib.lib().b2fhe_cleanup())
    _devices = []


def Synchronize():
    """cufhe::Synchronize (include/cufhe_gpu.cuh:68-74)."""
    _lib.check(_lib.lib().b2fhe_synchronize())


class Stream:
    """cufhe::Stream (include/cufhe_gpu.cuh:152-185): a CUDA stream bound to a device;
    default construction round-robins over the configured GPUs."""

    def __init__(self, device_id=None):
        global _stream_count
        if device_id is None:
            devs = _devices or [0]
            device_id = devs[_stream_count % len(devs)]
            _stream_count += 1
        self._device_id = int(device_id)
        self._st = None

    def Create(self):
        self._st = torch.cuda.Stream(device=self._device_id)
        return self

    def Destroy(self):
        self._st = None

    def st(self):
        return 0 if self._st is None else self._st.cuda_stream

    def device_id(self):
        return self._device_id

    def torch_stream(self):
        return self._st if self._st is not None else torch.cuda.default_stream(self._device_id)


def StreamQuery(st):
    """cufhe::StreamQuery (src/cufhe_gates_gpu.cu:55-65)."""
    return st.torch_stream().query()


class Ctxt:
    """A batch of TLWE ciphertexts: pinned host array + one device buffer
    (cufhe::Ctxt<P>, include/cufhe_gpu.cuh:102-121; P = lvl0param / lvl1param <-> level)."""

    def __init__(self, count, level=0, device=0):
        self.count, self.level, self.device = int(count), int(level), int(device)
        shape = (self.count, _words(level))
        self._host_t = torch.empty(shape, dtype=_torch_dtype(level)).pin_memory()
        self.tlwehost = self._host_t.numpy().view(_np_dtype(level))
        self.tlwedevice = torch.empty(shape, dtype=_torch_dtype(level), device=torch.device("cuda", self.device))

    def set(self, array):
        self.tlwehost[...] = np.asarray(array, dtype=_np_dtype(self.level)).reshape(self.tlwehost.shape)
        return self

    def ptr(self):
        return self.tlwedevice.data_ptr()


def CtxtCopyH2D(c, st):
    with torch.cuda.stream(st.torch_stream()):
        c.tlwedevice.copy_(c._host_t, non_blocking=True)


def CtxtCopyD2H(c, st):
    with torch.cuda.stream(st.torch_stream()):
        c._host_t.copy_(c.tlwedevice, non_blocking=True)


def _gate_device(op, out, in0, in1, in2, st):
    for c in (in0, in1, in2):
        if c is not None and (c.level != out.level or c.count != out.count or c.device != out.device):
            raise ValueError("operand batches must match the output's level, count and device")
    if st.device_id() != out.device:
        raise ValueError("stream and ciphertexts live on different devices")
    _lib.check(_lib.lib().b2fhe_gate_batch(
        out.device, st.st(), OP[op], out.level, out.ptr(), in0.ptr(),
        in1.ptr() if in1 is not None else None, in2.ptr() if in2 is not None else None, out.count))


def _gate_copying(op, out, in0, in1, in2, st):
    # the copying API: H2D inputs, gate, D2H output (src/cufhe_gates_gpu.cu:148-158)
    for c in (in0, in1, in2):
        if c is not None:
            CtxtCopyH2D(c, st)
    _gate_device(op, out, in0, in1, in2, st)
    CtxtCopyD2H(out, st)


def _make2(op):
    def g(out, in0, in1, st):
        _gate_device(op, out, in0, in1, None, st)

    def f(out, in0, in1, st):
        _gate_copying(op, out, in0, in1, None, st)
    return f, g


Nand, gNand = _make2("NAND")
Or, gOr = _make2("OR")
OrYN, gOrYN = _make2("ORYN")
OrNY, gOrNY = _make2("ORNY")
And, gAnd = _make2("AND")
AndYN, gAndYN = _make2("ANDYN")
AndNY, gAndNY = _make2("ANDNY")
Nor, gNor = _make2("NOR")
Xor, gXor = _make2("XOR")
Xnor, gXnor = _make2("XNOR")


def Mux(out, inc, in1, in0, st):
    _gate_copying("MUX", out, inc, in1, in0, st)


def gMux(out, inc, in1, in0, st):
    _gate_device("MUX", out, inc, in1, in0, st)


def NMux(out, inc, in1, in0, st):
    _gate_copying("NMUX", out, inc, in1, in0, st)


def gNMux(out, inc, in1, in0, st):
    _gate_device("NMUX", out, inc, in1, in0, st)


def Not(out, in0, st):
    _gate_copying("NOT", out, in0, None, None, st)


def gNot(out, in0, st):
    _gate_device("NOT", out, in0, None, None, st)


def Copy(out, in0, st):
    _gate_copying("COPY", out, in0, None, None, st)


def gCopy(out, in0, st):
    _gate_device("COPY", out, in0, None, None, st)


def CopyOnHost(out, in0):
    out.tlwehost[...] = in0.tlwehost


def gGateMixed(ops, out, in0, in1, in2, st):
    """Per-gate op codes over one batch (BASELINE config 4): ops is a uint8 array of
    len(out) op codes; two-input gates read (in0,in1), Mux/NMux read (inc=in0,in1,in2),
    Not/Copy read in0."""
    ops = np.ascontiguousarray(ops, dtype=np.uint8)
    if ops.size != out.count:
        raise ValueError("one op code per gate")
    _lib.check(_lib.lib().b2fhe_gate_batch_mixed(
        out.device, st.st(), ops.ctypes.data, out.level, out.ptr(), in0.ptr(), in1.ptr(), in2.ptr(), out.count))


class cuFHETRLWElvl1:
    """A batch of TRLWE<lvl1param> ([a(X)][b(X)], 2 x 1024 u32): pinned host array + device
    buffer (cufhe::cuFHETRLWElvl1, include/cufhe_gpu.cuh:123-134)."""

    def __init__(self, count, device=0):
        self.count, self.device = int(count), int(device)
        self._host_t = torch.empty((self.count, 2 * 1024), dtype=torch.int32).pin_memory()
        self.trlwehost = self._host_t.numpy().view(np.uint32)
        self.trlwedevice = torch.empty((self.count, 2 * 1024), dtype=torch.int32, device=torch.device("cuda", self.device))

    def set(self, array):
        self.trlwehost[...] = np.asarray(array, dtype=np.uint32).reshape(self.trlwehost.shape)
        return self

    def ptr(self):
        return self.trlwedevice.data_ptr()


def gBootstrap(out, in0, mu, st):
    """cufhe::Bootstrap (src/bootstrap_gpu.cu:782-789): lvl0 -> lvl0 bootstrapping with a
    run-time test-vector constant mu, on device-resident batches."""
    _lib.check(_lib.lib().b2fhe_bootstrap_batch(out.device, st.st(), out.ptr(), in0.ptr(), mu, out.count))


def gGateBootstrappingTLWE2TRLWElvl01NTT(out, in0, st, mu=1 << 29):
    """cufhe::gGateBootstrappingTLWE2TRLWElvl01NTT (src/cufhe_gates_gpu.cu:101-108)."""
    _lib.check(_lib.lib().b2fhe_bootstrap_tlwe2trlwe_batch(out.device, st.st(), out.ptr(), in0.ptr(), mu, out.count))


def GateBootstrappingTLWE2TRLWElvl01NTT(out, in0, st, mu=1 << 29):
    """copying variant (src/cufhe_gates_gpu.cu:88-99)."""
    CtxtCopyH2D(in0, st)
    gGateBootstrappingTLWE2TRLWElvl01NTT(out, in0, st, mu)
    with torch.cuda.stream(st.torch_stream()):
        out._host_t.copy_(out.trlwedevice, non_blocking=True)


def gSampleExtractAndKeySwitch(out, in0, st):
    """cufhe::gSampleExtractAndKeySwitch on a device-resident TRLWE batch."""
    _lib.check(_lib.lib().b2fhe_sample_extract_keyswitch_batch(out.device, st.st(), out.ptr(), in0.ptr(), out.count))


def SampleExtractAndKeySwitch(out, in0, st):
    """cufhe::SampleExtractAndKeySwitch (src/cufhe_gates_gpu.cu:128-137)."""
    with torch.cuda.stream(st.torch_stream()):
        in0.trlwedevice.copy_(in0._host_t, non_blocking=True)
    gSampleExtractAndKeySwitch(out, in0, st)
    CtxtCopyD2H(out, st)


def gRefresh(out, in0, st, mu=1 << 29):
    """cufhe::gRefresh with the reference's body-term bug fixed (include/b2fhe.h)."""
    _lib.check(_lib.lib().b2fhe_refresh_batch(out.device, st.st(), out.ptr(), in0.ptr(), mu, out.count))


def Refresh(out, in0, st, mu=1 << 29):
    """cufhe::Refresh (src/cufhe_gates_gpu.cu:110-119)."""
    with torch.cuda.stream(st.torch_stream()):
        in0.trlwedevice.copy_(in0._host_t, non_blocking=True)
    gRefresh(out, in0, st, mu)
    with torch.cuda.stream(st.torch_stream()):
        out._host_t.copy_(out.trlwedevice, non_blocking=True)


class cuFHETRGSWNTTlvl1:
    """A batch of NTT-domain TRGSWs in this engine's format (B2FHE_TRGSW_NTT_BYTES each);
    produced only by TRGSW2NTT (cufhe::cuFHETRGSWNTTlvl1, include/cufhe_gpu.cuh:136-146)."""

    def __init__(self, count, device=0):
        self.count, self.device = int(count), int(device)
        self.trgswdevice = torch.empty((self.count, 16384), dtype=torch.int32, device=torch.device("cuda", self.device))

    def ptr(self):
        return self.trgswdevice.data_ptr()


def TRGSW2NTT(trgswntt, trgsw_raw, st):
    """cufhe::TRGSW2NTT (src/bootstrap_gpu.cu:75-95): raw TRGSW<lvl1param> [count][4][2][1024]
    u32 (host numpy) -> NTT domain on the device."""
    raw = np.ascontiguousarray(trgsw_raw, dtype=np.uint32).reshape(trgswntt.count, 8192)
    with torch.cuda.stream(st.torch_stream()):
        d = torch.from_numpy(raw.view(np.int32)).to(trgswntt.trgswdevice.device, non_blocking=False)
        _lib.check(_lib.lib().b2fhe_trgsw2ntt_batch(trgswntt.device, st.st(), trgswntt.ptr(), d.data_ptr(), trgswntt.count))
        d.record_stream(st.torch_stream())


def gCMUXNTT(res, cs, c1, c0, st):
    """cufhe::CMUXNTT on device-resident operands: res = c0 + (c1 - c0) [.] cs.  One TRGSW per
    item if cs.count == res.count, one shared TRGSW if cs.count == 1."""
    if cs.count not in (1, res.count):
        raise ValueError("cs must hold one TRGSW or one per item")
    _lib.check(_lib.lib().b2fhe_cmux_ntt_batch(res.device, st.st(), res.ptr(), cs.ptr(), c1.ptr(), c0.ptr(), res.count,
                                               1 if cs.count == res.count and res.count > 1 else 0))


def CMUXNTT(res, cs, c1, c0, st):
    """copying variant (src/cufhe_gates_gpu.cu:69-85)."""
    with torch.cuda.stream(st.torch_stream()):
        c1.trlwedevice.copy_(c1._host_t, non_blocking=True)
        c0.trlwedevice.copy_(c0._host_t, non_blocking=True)
    gCMUXNTT(res, cs, c1, c0, st)
    with torch.cuda.stream(st.torch_stream()):
        res._host_t.copy_(res.trlwedevice, non_blocking=True)


def partition(count, world, rank):
    """Static contiguous partition of a gate batch across GPUs:
    returns (begin, end) of this rank's slice; sizes differ by at most one."""
    base, rem = divmod(int(count), int(world))
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)
