// cufhe_ref_gpu.cu -- thin C wrapper over the REFERENCE cuFHE GPU library itself, compiled
// from the sources where they lie under /root/reference (oracle/Makefile target
// _ref/libcufhe_ref.so; arch sm_100 instead of the reference's pinned sm_80, nothing else
// changed).  TEST INFRASTRUCTURE ONLY: tests/test_gpu_reference_cufhe.py runs it on the
// B200 and byte-compares its ciphertexts with the CPU oracle and with our engine on the
// same keys and inputs -- the "bit-exact against the reference cuFHE build" gate of
// BASELINE.json.  It only calls the reference's public API
// (include/cufhe_gpu.cuh: SetGPUNum, Initialize(EvalKey), Stream, Ctxt<P>, Nand<P> ..., Synchronize).
#include <cstring>
#include <memory>
#include <vector>
#include <include/cufhe_gpu.cuh>

using namespace TFHEpp;
static std::unique_ptr<EvalKey> g_ek;

template <class P>
static int run_gate(int op, void* out, const void* in0, const void* in1, const void* in2, int count)
{
    using namespace cufhe;
    const size_t sz = sizeof(TLWE<P>);
    Stream st;
    st.Create();
    // one Ctxt set per gate, as the reference's tests do (test/test_util.h:20-60)
    std::vector<Ctxt<P>> a(count), b(count), c(count), r(count);
    for (int i = 0; i < count; i++) {
        std::memcpy(a[i].tlwehost.data(), (const char*)in0 + i * sz, sz);
        if (in1) std::memcpy(b[i].tlwehost.data(), (const char*)in1 + i * sz, sz);
        if (in2) std::memcpy(c[i].tlwehost.data(), (const char*)in2 + i * sz, sz);
    }
    for (int i = 0; i < count; i++) {
        switch (op) {      // op codes of include/b2fhe.h
            case 0: Nand<P>(r[i], a[i], b[i], st); break;
            case 1: Or<P>(r[i], a[i], b[i], st); break;
            case 2: OrYN<P>(r[i], a[i], b[i], st); break;
            case 3: OrNY<P>(r[i], a[i], b[i], st); break;
            case 4: And<P>(r[i], a[i], b[i], st); break;
            case 5: AndYN<P>(r[i], a[i], b[i], st); break;
            case 6: AndNY<P>(r[i], a[i], b[i], st); break;
            case 7: Nor<P>(r[i], a[i], b[i], st); break;
            case 8: Xor<P>(r[i], a[i], b[i], st); break;
            case 9: Xnor<P>(r[i], a[i], b[i], st); break;
            case 10: Mux<P>(r[i], a[i], b[i], c[i], st); break;     // Mux(out, inc, in1, in0)
            case 11: NMux<P>(r[i], a[i], b[i], c[i], st); break;
            case 12: Not<P>(r[i], a[i], st); break;
            case 13: Copy<P>(r[i], a[i], st); break;
            default: return -1;
        }
    }
    Synchronize();
    for (int i = 0; i < count; i++) std::memcpy((char*)out + i * sz, r[i].tlwehost.data(), sz);
    st.Destroy();
    return 0;
}

extern "C" {
// bk_raw / ksk_raw: TFHEpp raw arrays (same bytes our b2fhe_init takes)
int cufheref_init(const void* bk_raw, const void* ksk_raw)
{
    g_ek = std::make_unique<EvalKey>();
    g_ek->get<BootstrappingKey<lvl01param>>() = std::make_unique_for_overwrite<BootstrappingKey<lvl01param>>();
    g_ek->get<KeySwitchingKey<lvl10param>>() = std::make_unique_for_overwrite<KeySwitchingKey<lvl10param>>();
    std::memcpy(&g_ek->getbk<lvl01param>(), bk_raw, sizeof(BootstrappingKey<lvl01param>));
    std::memcpy(&g_ek->getiksk<lvl10param>(), ksk_raw, sizeof(KeySwitchingKey<lvl10param>));
    cufhe::SetGPUNum(1);
    cufhe::Initialize(*g_ek);
    return 0;
}
int cufheref_gate(int op, int level, void* out, const void* in0, const void* in1, const void* in2, int count)
{
    return level == 0 ? run_gate<lvl0param>(op, out, in0, in1, in2, count)
                      : run_gate<lvl1param>(op, out, in0, in1, in2, count);
}
// the reference's own throughput method (test/test_gate_gpu.cc:41-83, test/test_util.h:29-72):
// `count` lvl0 NANDs issued round-robin over `nstreams` streams, H2D + kernel + D2H per gate;
// returns seconds between first issue and Synchronize().
double cufheref_bench_nand(int count, int nstreams)
{
    using namespace cufhe;
    using P = lvl0param;
    std::vector<Stream> st(nstreams);
    for (auto& s : st) s.Create();
    std::vector<Ctxt<P>> a(count), b(count), r(count);
    for (int i = 0; i < count; i++)
        for (size_t k = 0; k < a[i].tlwehost.size(); k++) {
            a[i].tlwehost[k] = (P::T)(i * 7919u + k * 104729u);
            b[i].tlwehost[k] = (P::T)(i * 15485863u + k * 32452843u);
        }
    for (int i = 0; i < nstreams && i < count; i++) Nand<P>(r[i], a[i], b[i], st[i]);   // warm-up
    Synchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, 0);
    for (int i = 0; i < count; i++) Nand<P>(r[i], a[i], b[i], st[i % nstreams]);
    Synchronize();
    cudaEventRecord(e1, 0);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    for (auto& s : st) s.Destroy();
    return ms * 1e-3;
}
void cufheref_cleanup(void) { cufhe::CleanUp(); g_ek.reset(); }
}
