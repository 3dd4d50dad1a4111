// b2fhe_noalloc.inl -- a gate entry point that allocates NOTHING (included by b2fhe_api.cu, so the
// array below is a file-scope object of that one translation unit).
//
// No counterpart in the reference: added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  b2fhe_gate_batch takes its stage scratch (the lvl1 TLWEs between blind rotation and key
// switch, 4100 B per gate) from the stream-ordered pool; this variant keeps that scratch in a fixed
// device array, so a call is exactly two kernel launches (three for Mux/NMux is NOT offered here).
//
// Sharing rules: every call on a device uses the SAME array.  Calls are safe back to back on one
// stream (stream order); calls that overlap on the GPU (different streams, no dependency) RACE.
// Nothing in the array has to start at zero: the blind rotation writes all 1025 words of every slot
// the key switch then reads, and slots beyond `count` are never read.

namespace {

// lvl1 TLWE scratch: B2FHE_NOALLOC_MAX_GATES x 1025 words = 8,396,800 B, held by every GPU that loads
// the library until it exits
static __device__ __align__(256) uint32_t g_noalloc_scratch[(size_t)B2FHE_NOALLOC_MAX_GATES * kLvl1Words];
uint32_t* g_noalloc_ptr[kMaxDev] = {};

// set-up, once per device, with that device current: in host code `g_noalloc_scratch` is only a
// stand-in for the symbol -- the device address comes from cudaGetSymbolAddress
int noalloc_setup(int device)
{
    if (g_noalloc_ptr[device]) return 0;
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    REQUIRE(p != nullptr, B2FHE_E_NO_DEVICE);
    g_noalloc_ptr[device] = static_cast<uint32_t*>(p);
    return 0;
}

}  // namespace

extern "C" {

size_t b2fhe_noalloc_max_gates(void) { return B2FHE_NOALLOC_MAX_GATES; }

int b2fhe_gate_batch_noalloc(int device, b2fhe_stream_t stream, int op, uint16_t* out_lvl0, const uint16_t* in0_lvl0,
                             const uint16_t* in1_lvl0, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(op >= 0 && op < B2FHE_MUX, B2FHE_E_BAD_ARGUMENT);                 // the ten two-input gates
    if (count > B2FHE_NOALLOC_MAX_GATES) {
        char msg[160];
        snprintf(msg, sizeof msg, "b2fhe_gate_batch_noalloc: count %zu exceeds the fixed scratch of %d gates", count,
                 B2FHE_NOALLOC_MAX_GATES);
        t_err = msg;
        return B2FHE_E_TOO_LARGE;
    }
    if (count == 0) return 0;
    REQUIRE(out_lvl0 && in0_lvl0 && in1_lvl0, B2FHE_E_BAD_ARGUMENT);
    uint32_t* t1 = g_noalloc_ptr[device];
    REQUIRE(t1 != nullptr, B2FHE_E_NOT_INITIALIZED);
    cudaStream_t st = (cudaStream_t)stream;
    int ca, cb, so;
    gate_coeffs(op, ca, cb, so);
    // BR(pre-add) -> SE -> IKS, as uniform_gate's lvl0 branch (src/bootstrap_gpu.cu:402-421)
    BrParams B{};
    B.in0 = in0_lvl0; B.in1 = in1_lvl0; B.out = t1; B.count = (int)count;
    B.ca = ca; B.cb = cb; B.off = so * (int)kMu0; B.mu = kMu1;
    if (int r = launch_br(*c, B, st)) return r;
    KsParams K{};
    K.A = t1; K.out = out_lvl0; K.count = (int)count; K.ca = 1; K.cb = 0; K.off = 0u;
    return launch_ks(*c, K, st);
}

}  // extern "C"
