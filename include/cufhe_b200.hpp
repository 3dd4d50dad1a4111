// cufhe_b200.hpp -- header-only C++ shim: cuFHE's public API (namespace cufhe,
// reference include/cufhe_gpu.cuh:42-315 and src/cufhe_gates_gpu.cu) implemented as
// thin calls into the C ABI of libb2fhe.so (include/b2fhe.h).  Same names, argument
// order and semantics, so code written against cuFHE's gate API recompiles unchanged:
//
//   SetGPUNum, Initialize(EvalKey), CleanUp, Synchronize, StreamQuery, Stream,
//   Ctxt<P>, CtxtCopyH2D/D2H, And ... Xnor, Mux, NMux, Not, Copy, CopyOnHost and the
//   device-resident g* twins, for P = TFHEpp::lvl0param / lvl1param.
//
// Added beside it (not in the reference): CtxtBatch<P> + <Gate>Batch(...) -- one call,
// one batched launch -- which is how the engine reaches its throughput.
// Error convention of the reference is kept: a failing call prints the message and
// exit(-1)s (include/details/error_gpu.cuh:30-66).
#pragma once
#include <cuda_runtime.h>

#include <array>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "b2fhe.h"
#ifdef B2FHE_WITH_TFHEPP
#include <tfhe++.hpp>
#else
#include "tfhepp_compat.hpp"
#endif

namespace cufhe {

// the reference's public globals (include/cufhe_gpu.cuh:44-46, src/cufhe_gates_gpu.cu:34-36): C++17 inline
// variables, so this header-only shim still gives ONE cufhe::_gpuNum / cufhe::streamCount per program
inline int _gpuNum = 1;
inline int streamCount = 0;

inline void _b2check(int rc, const char* file, int line)
{
    if (rc != 0) {
        fprintf(stderr, "cufhe(b2fhe) failed at %s:%i : %s\n", file, line, b2fhe_last_error_string());
        exit(-1);
    }
}
#define B2CHECK(x) ::cufhe::_b2check((x), __FILE__, __LINE__)

template <class P> constexpr int _level() { return sizeof(typename P::T) == 2 ? 0 : 1; }

inline void SetGPUNum(int gpuNum) { _gpuNum = gpuNum; }
// Initialize NTT handlers only (include/cufhe_gpu.cuh:56-57, src/cufhe_gates_gpu.cu:40): twiddle tables and
// kernel attributes on every device, no keys
inline void Initialize() { B2CHECK(b2fhe_setup_devices(_gpuNum)); }
inline void Initialize(const TFHEpp::EvalKey& ek)
{
    B2CHECK(b2fhe_init(_gpuNum, reinterpret_cast<const uint32_t*>(&ek.getbk<TFHEpp::lvl01param>()),
                       reinterpret_cast<const uint16_t*>(&ek.getiksk<TFHEpp::lvl10param>())));
}
// Not in the reference: Initialize from the bytes of a serialised TFHEpp::EvalKey (cereal
// portable_binary, TFHEpp/include/cloudkey.hpp:103-107 -- e.g. the tutorial's cloud.key) without
// linking TFHEpp's loader, and from / to a cache of the converted keys (skips the conversion).
inline void InitializeFromArchive(const void* blob, size_t len) { B2CHECK(b2fhe_init_from_evalkey(_gpuNum, blob, len)); }
inline void InitializeFromKeyCache(const void* blob, size_t len) { B2CHECK(b2fhe_init_from_key_cache(_gpuNum, blob, len)); }
inline std::vector<uint8_t> ExportKeyCache(int device = 0)
{
    std::vector<uint8_t> blob(b2fhe_key_cache_size());
    B2CHECK(b2fhe_key_cache_export(device, blob.data(), blob.size(), nullptr));
    return blob;
}
inline void CleanUp() { B2CHECK(b2fhe_cleanup()); }
// Synchronize(): runs the queued per-gate calls (see SetDeferred) and waits for every device
inline void Synchronize() { B2CHECK(b2fhe_synchronize()); }
// Per-gate calls (Nand(out, in0, in1, st) ... on single Ctxt<P>) are QUEUED and executed at
// Synchronize()/StreamQuery() as batched launches levelised by their data dependencies --
// one kernel per gate, the reference's execution model, starves this engine's batch-major
// kernels.  Results are defined exactly when the reference defines them: after
// Synchronize() or a successful StreamQuery().  SetDeferred(false) restores immediate
// one-launch-per-call execution (lowest latency for a single isolated gate).
inline bool& _deferredRef() { static bool on = true; return on; }
inline void SetDeferred(bool on)
{
    if (!on) Synchronize();
    _deferredRef() = on;
}

class Stream {
public:
    Stream() : st_(0), device_id_(streamCount++ % _gpuNum) {}
    explicit Stream(int device_id) : st_(0), device_id_(device_id) { streamCount++; }
    void Create()
    {
        cudaSetDevice(device_id_);
        cudaStreamCreateWithFlags(&st_, cudaStreamNonBlocking);
    }
    void Destroy()
    {
        cudaSetDevice(device_id_);
        cudaStreamDestroy(st_);
        st_ = 0;
    }
    cudaStream_t st() const { return st_; }
    int device_id() const { return device_id_; }

private:
    cudaStream_t st_;
    int device_id_;
};
inline bool StreamQuery(Stream st)
{
    int busy = 0;
    B2CHECK(b2fhe_defer_query_stream(st.device_id(), (b2fhe_stream_t)st.st(), &busy));      // this stream's queued gates start now
    cudaSetDevice(st.device_id());
    return !busy && cudaStreamQuery(st.st()) == cudaSuccess;
}

// a batch of `count` ciphertexts: pinned host array + one device buffer per GPU
template <class P>
struct CtxtBatch {
    explicit CtxtBatch(size_t count) : count(count)
    {
        const size_t bytes = count * sizeof(TFHEpp::TLWE<P>);
        cudaMallocHost((void**)&tlwehost, bytes);
        tlwedevices.resize(_gpuNum);
        for (int i = 0; i < _gpuNum; i++) {
            cudaSetDevice(i);
            cudaMalloc((void**)&tlwedevices[i], bytes);
        }
    }
    ~CtxtBatch()
    {
        cudaFreeHost(tlwehost);
        for (size_t i = 0; i < tlwedevices.size(); i++) {
            cudaSetDevice((int)i);
            cudaFree(tlwedevices[i]);
        }
    }
    CtxtBatch(const CtxtBatch&) = delete;
    CtxtBatch& operator=(const CtxtBatch&) = delete;
    size_t bytes() const { return count * sizeof(TFHEpp::TLWE<P>); }
    size_t count;
    TFHEpp::TLWE<P>* tlwehost;
    std::vector<typename P::T*> tlwedevices;
};

// cufhe::Ctxt<P> (include/cufhe_gpu.cuh:102-121): one ciphertext
template <class P>
struct Ctxt {
    Ctxt()
    {
        cudaHostRegister(tlwehost.data(), sizeof(tlwehost), cudaHostRegisterDefault);
        tlwedevices.resize(_gpuNum);
        for (int i = 0; i < _gpuNum; i++) {
            cudaSetDevice(i);
            cudaMalloc((void**)&tlwedevices[i], sizeof(tlwehost));
        }
    }
    ~Ctxt()
    {
        for (int i = 0; i < (int)tlwedevices.size(); i++)
            if (b2fhe_defer_pending(i) > 0) { b2fhe_synchronize(); break; }      // queued gates may use this buffer
        cudaHostUnregister(tlwehost.data());
        for (size_t i = 0; i < tlwedevices.size(); i++) {
            cudaSetDevice((int)i);
            cudaFree(tlwedevices[i]);
        }
    }
    Ctxt(const Ctxt&) = delete;
    Ctxt& operator=(const Ctxt&) = delete;
    alignas(64) TFHEpp::TLWE<P> tlwehost;
    std::vector<typename P::T*> tlwedevices;
};

template <class P> inline void CtxtCopyH2D(Ctxt<P>& c, Stream st)
{
    // stream order: a copy into a buffer that a queued gate still reads or writes runs after it
    B2CHECK(b2fhe_defer_fence_buffer(st.device_id(), (b2fhe_stream_t)st.st(), c.tlwedevices[st.device_id()]));
    cudaSetDevice(st.device_id());
    cudaMemcpyAsync(c.tlwedevices[st.device_id()], c.tlwehost.data(), sizeof(c.tlwehost), cudaMemcpyHostToDevice, st.st());
}
template <class P> inline void CtxtCopyD2H(Ctxt<P>& c, Stream st)
{
    B2CHECK(b2fhe_defer_fence(st.device_id(), (b2fhe_stream_t)st.st()));
    cudaSetDevice(st.device_id());
    cudaMemcpyAsync(c.tlwehost.data(), c.tlwedevices[st.device_id()], sizeof(c.tlwehost), cudaMemcpyDeviceToHost, st.st());
}
template <class P> inline void CtxtCopyH2D(CtxtBatch<P>& c, Stream st)
{
    B2CHECK(b2fhe_defer_fence_buffer(st.device_id(), (b2fhe_stream_t)st.st(), c.tlwedevices[st.device_id()]));
    cudaSetDevice(st.device_id());
    cudaMemcpyAsync(c.tlwedevices[st.device_id()], c.tlwehost, c.bytes(), cudaMemcpyHostToDevice, st.st());
}
template <class P> inline void CtxtCopyD2H(CtxtBatch<P>& c, Stream st)
{
    B2CHECK(b2fhe_defer_fence(st.device_id(), (b2fhe_stream_t)st.st()));
    cudaSetDevice(st.device_id());
    cudaMemcpyAsync(c.tlwehost, c.tlwedevices[st.device_id()], c.bytes(), cudaMemcpyDeviceToHost, st.st());
}
template <class P> inline void CopyOnHost(Ctxt<P>& out, Ctxt<P>& in) { out.tlwehost = in.tlwehost; }

// ---- gates ---------------------------------------------------------------------------
template <class P> inline void* _hostptr(Ctxt<P>& c) { return c.tlwehost.data(); }
template <class P> inline void* _hostptr(CtxtBatch<P>& c) { return c.tlwehost; }
template <class P, class C>
inline void _gate(int op, C& out, C* in0, C* in1, C* in2, size_t count, Stream st, bool copy)
{
    const int d = st.device_id();
    if (count == 1 && _deferredRef()) {
        // legacy one-gate call: operand uploads, the gate and the copy-back are all queued and
        // run as batched transfers + launches when the queue is flushed
        if (copy) {
            B2CHECK(b2fhe_defer_h2d(d, in0->tlwedevices[d], _hostptr<P>(*in0), sizeof(TFHEpp::TLWE<P>)));
            if (in1) B2CHECK(b2fhe_defer_h2d(d, in1->tlwedevices[d], _hostptr<P>(*in1), sizeof(TFHEpp::TLWE<P>)));
            if (in2) B2CHECK(b2fhe_defer_h2d(d, in2->tlwedevices[d], _hostptr<P>(*in2), sizeof(TFHEpp::TLWE<P>)));
        }
        B2CHECK(b2fhe_defer_gate(d, (b2fhe_stream_t)st.st(), op, _level<P>(), out.tlwedevices[d], in0->tlwedevices[d],
                                 in1 ? in1->tlwedevices[d] : nullptr, in2 ? in2->tlwedevices[d] : nullptr));
        if (copy) B2CHECK(b2fhe_defer_d2h(d, _hostptr<P>(out), out.tlwedevices[d], sizeof(TFHEpp::TLWE<P>)));
        return;
    }
    B2CHECK(b2fhe_defer_fence(d, (b2fhe_stream_t)st.st()));     // immediate call: after anything queued
    if (copy) {
        CtxtCopyH2D<P>(*in0, st);
        if (in1) CtxtCopyH2D<P>(*in1, st);
        if (in2) CtxtCopyH2D<P>(*in2, st);
    }
    B2CHECK(b2fhe_gate_batch(d, (b2fhe_stream_t)st.st(), op, _level<P>(), out.tlwedevices[d], in0->tlwedevices[d],
                             in1 ? in1->tlwedevices[d] : nullptr, in2 ? in2->tlwedevices[d] : nullptr, count));
    if (copy) CtxtCopyD2H<P>(out, st);
}

#define B2_GATE2(Name, OP)                                                                                         \
    template <class P> inline void Name(Ctxt<P>& out, Ctxt<P>& in0, Ctxt<P>& in1, Stream st)                        \
    { _gate<P>(OP, out, &in0, &in1, (Ctxt<P>*)nullptr, 1, st, true); }                                              \
    template <class P> inline void g##Name(Ctxt<P>& out, Ctxt<P>& in0, Ctxt<P>& in1, Stream st)                     \
    { _gate<P>(OP, out, &in0, &in1, (Ctxt<P>*)nullptr, 1, st, false); }                                             \
    template <class P> inline void Name##Batch(CtxtBatch<P>& out, CtxtBatch<P>& in0, CtxtBatch<P>& in1, Stream st)  \
    { _gate<P>(OP, out, &in0, &in1, (CtxtBatch<P>*)nullptr, out.count, st, true); }                                 \
    template <class P> inline void g##Name##Batch(CtxtBatch<P>& out, CtxtBatch<P>& in0, CtxtBatch<P>& in1, Stream st) \
    { _gate<P>(OP, out, &in0, &in1, (CtxtBatch<P>*)nullptr, out.count, st, false); }
B2_GATE2(Nand, B2FHE_NAND)
B2_GATE2(Or, B2FHE_OR)
B2_GATE2(OrYN, B2FHE_ORYN)
B2_GATE2(OrNY, B2FHE_ORNY)
B2_GATE2(And, B2FHE_AND)
B2_GATE2(AndYN, B2FHE_ANDYN)
B2_GATE2(AndNY, B2FHE_ANDNY)
B2_GATE2(Nor, B2FHE_NOR)
B2_GATE2(Xor, B2FHE_XOR)
B2_GATE2(Xnor, B2FHE_XNOR)
#undef B2_GATE2

#define B2_GATE3(Name, OP)                                                                                          \
    template <class P> inline void Name(Ctxt<P>& out, Ctxt<P>& inc, Ctxt<P>& in1, Ctxt<P>& in0, Stream st)          \
    { _gate<P>(OP, out, &inc, &in1, &in0, 1, st, true); }                                                           \
    template <class P> inline void g##Name(Ctxt<P>& out, Ctxt<P>& inc, Ctxt<P>& in1, Ctxt<P>& in0, Stream st)       \
    { _gate<P>(OP, out, &inc, &in1, &in0, 1, st, false); }                                                          \
    template <class P> inline void Name##Batch(CtxtBatch<P>& out, CtxtBatch<P>& inc, CtxtBatch<P>& in1, CtxtBatch<P>& in0, Stream st) \
    { _gate<P>(OP, out, &inc, &in1, &in0, out.count, st, true); }                                                   \
    template <class P> inline void g##Name##Batch(CtxtBatch<P>& out, CtxtBatch<P>& inc, CtxtBatch<P>& in1, CtxtBatch<P>& in0, Stream st) \
    { _gate<P>(OP, out, &inc, &in1, &in0, out.count, st, false); }
B2_GATE3(Mux, B2FHE_MUX)
B2_GATE3(NMux, B2FHE_NMUX)
#undef B2_GATE3

#define B2_GATE1(Name, OP)                                                                                          \
    template <class P> inline void Name(Ctxt<P>& out, Ctxt<P>& in, Stream st)                                       \
    { _gate<P>(OP, out, &in, (Ctxt<P>*)nullptr, (Ctxt<P>*)nullptr, 1, st, true); }                                  \
    template <class P> inline void g##Name(Ctxt<P>& out, Ctxt<P>& in, Stream st)                                    \
    { _gate<P>(OP, out, &in, (Ctxt<P>*)nullptr, (Ctxt<P>*)nullptr, 1, st, false); }                                 \
    template <class P> inline void Name##Batch(CtxtBatch<P>& out, CtxtBatch<P>& in, Stream st)                      \
    { _gate<P>(OP, out, &in, (CtxtBatch<P>*)nullptr, (CtxtBatch<P>*)nullptr, out.count, st, true); }                \
    template <class P> inline void g##Name##Batch(CtxtBatch<P>& out, CtxtBatch<P>& in, Stream st)                   \
    { _gate<P>(OP, out, &in, (CtxtBatch<P>*)nullptr, (CtxtBatch<P>*)nullptr, out.count, st, false); }
B2_GATE1(Not, B2FHE_NOT)
B2_GATE1(Copy, B2FHE_COPY)
#undef B2_GATE1

// ---- TRLWE-level entry points (include/cufhe_gpu.cuh:123-146,209-216,282-285) -------------
struct cuFHETRLWElvl1 {
    cuFHETRLWElvl1()
    {
        cudaHostRegister(trlwehost.data(), sizeof(trlwehost), cudaHostRegisterDefault);
        trlwedevices.resize(_gpuNum);
        for (int i = 0; i < _gpuNum; i++) {
            cudaSetDevice(i);
            cudaMalloc((void**)&trlwedevices[i], sizeof(trlwehost));
        }
    }
    ~cuFHETRLWElvl1()
    {
        cudaHostUnregister(trlwehost.data());
        for (size_t i = 0; i < trlwedevices.size(); i++) {
            cudaSetDevice((int)i);
            cudaFree(trlwedevices[i]);
        }
    }
    cuFHETRLWElvl1(const cuFHETRLWElvl1&) = delete;
    cuFHETRLWElvl1& operator=(const cuFHETRLWElvl1&) = delete;
    TFHEpp::TRLWE<TFHEpp::lvl1param> trlwehost;
    std::vector<TFHEpp::lvl1param::T*> trlwedevices;
};
inline void _trlweH2D(const cuFHETRLWElvl1& c, Stream st)
{
    cudaMemcpyAsync(c.trlwedevices[st.device_id()], c.trlwehost.data(), sizeof(c.trlwehost), cudaMemcpyHostToDevice, st.st());
}
inline void _trlweD2H(cuFHETRLWElvl1& c, Stream st)
{
    cudaMemcpyAsync(c.trlwehost.data(), c.trlwedevices[st.device_id()], sizeof(c.trlwehost), cudaMemcpyDeviceToHost, st.st());
}
// Bootstrap (include/bootstrap_gpu.cuh:64-65): lvl0 -> lvl0 with a run-time test-vector constant
inline void gBootstrap(Ctxt<TFHEpp::lvl0param>& out, Ctxt<TFHEpp::lvl0param>& in, TFHEpp::lvl1param::T mu, Stream st)
{
    const int d = st.device_id();
    cudaSetDevice(d);
    B2CHECK(b2fhe_defer_fence(d, (b2fhe_stream_t)st.st()));
    B2CHECK(b2fhe_bootstrap_batch(d, (b2fhe_stream_t)st.st(), out.tlwedevices[d], in.tlwedevices[d], mu, 1));
}
inline void gGateBootstrappingTLWE2TRLWElvl01NTT(cuFHETRLWElvl1& out, Ctxt<TFHEpp::lvl0param>& in, Stream st)
{
    const int d = st.device_id();
    cudaSetDevice(d);
    B2CHECK(b2fhe_defer_fence(d, (b2fhe_stream_t)st.st()));
    B2CHECK(b2fhe_bootstrap_tlwe2trlwe_batch(d, (b2fhe_stream_t)st.st(), out.trlwedevices[d], in.tlwedevices[d],
                                             TFHEpp::lvl1param::μ, 1));
}
inline void GateBootstrappingTLWE2TRLWElvl01NTT(cuFHETRLWElvl1& out, Ctxt<TFHEpp::lvl0param>& in, Stream st)
{
    CtxtCopyH2D<TFHEpp::lvl0param>(in, st);
    gGateBootstrappingTLWE2TRLWElvl01NTT(out, in, st);
    _trlweD2H(out, st);
}
inline void gSampleExtractAndKeySwitch(Ctxt<TFHEpp::lvl0param>& out, const cuFHETRLWElvl1& in, Stream st)
{
    const int d = st.device_id();
    cudaSetDevice(d);
    _trlweH2D(in, st);      // the reference's g-variant copies too (src/cufhe_gates_gpu.cu:139-146)
    B2CHECK(b2fhe_defer_fence(d, (b2fhe_stream_t)st.st()));
    B2CHECK(b2fhe_sample_extract_keyswitch_batch(d, (b2fhe_stream_t)st.st(), out.tlwedevices[d], in.trlwedevices[d], 1));
}
inline void SampleExtractAndKeySwitch(Ctxt<TFHEpp::lvl0param>& out, const cuFHETRLWElvl1& in, Stream st)
{
    gSampleExtractAndKeySwitch(out, in, st);
    CtxtCopyD2H<TFHEpp::lvl0param>(out, st);
}
inline void gRefresh(cuFHETRLWElvl1& out, cuFHETRLWElvl1& in, Stream st)
{
    const int d = st.device_id();
    cudaSetDevice(d);
    B2CHECK(b2fhe_defer_fence(d, (b2fhe_stream_t)st.st()));
    B2CHECK(b2fhe_refresh_batch(d, (b2fhe_stream_t)st.st(), out.trlwedevices[d], in.trlwedevices[d], TFHEpp::lvl1param::μ, 1));
}
inline void Refresh(cuFHETRLWElvl1& out, cuFHETRLWElvl1& in, Stream st)
{
    cudaSetDevice(st.device_id());
    _trlweH2D(in, st);
    gRefresh(out, in, st);
    _trlweD2H(out, st);
}

// ---- CMUX with a caller-supplied TRGSW (include/cufhe_gpu.cuh:136-146,209-216) -------------
// trgswhost holds this engine's NTT-domain form (B2FHE_TRGSW_NTT_BYTES, opaque) instead of
// TFHEpp::TRGSWNTT<lvl1param>: like the reference's, it is only ever produced by TRGSW2NTT.
struct cuFHETRGSWNTTlvl1 {
    cuFHETRGSWNTTlvl1()
    {
        cudaHostRegister(trgswhost.data(), sizeof(trgswhost), cudaHostRegisterDefault);
        trgswdevices.resize(_gpuNum);
        for (int i = 0; i < _gpuNum; i++) {
            cudaSetDevice(i);
            cudaMalloc((void**)&trgswdevices[i], sizeof(trgswhost));
        }
    }
    ~cuFHETRGSWNTTlvl1()
    {
        cudaHostUnregister(trgswhost.data());
        for (size_t i = 0; i < trgswdevices.size(); i++) {
            cudaSetDevice((int)i);
            cudaFree(trgswdevices[i]);
        }
    }
    cuFHETRGSWNTTlvl1(const cuFHETRGSWNTTlvl1&) = delete;
    cuFHETRGSWNTTlvl1& operator=(const cuFHETRGSWNTTlvl1&) = delete;
    std::array<uint32_t, B2FHE_TRGSW_NTT_BYTES / 4> trgswhost;
    std::vector<uint32_t*> trgswdevices;
};
// TRGSW2NTT (src/bootstrap_gpu.cu:75-95): raw TRGSW -> NTT domain on the device, result also
// copied back to trgswhost
inline void TRGSW2NTT(cuFHETRGSWNTTlvl1& trgswntt, const TFHEpp::TRGSW<TFHEpp::lvl1param>& trgsw, Stream st)
{
    const int d = st.device_id();
    cudaSetDevice(d);
    uint32_t* raw = nullptr;
    cudaMallocAsync((void**)&raw, sizeof(trgsw), st.st());
    cudaMemcpyAsync(raw, trgsw.data(), sizeof(trgsw), cudaMemcpyHostToDevice, st.st());
    B2CHECK(b2fhe_defer_fence(d, (b2fhe_stream_t)st.st()));
    B2CHECK(b2fhe_trgsw2ntt_batch(d, (b2fhe_stream_t)st.st(), trgswntt.trgswdevices[d], raw, 1));
    cudaFreeAsync(raw, st.st());
    cudaMemcpyAsync(trgswntt.trgswhost.data(), trgswntt.trgswdevices[d], sizeof(trgswntt.trgswhost), cudaMemcpyDeviceToHost, st.st());
}
inline void gCMUXNTT(cuFHETRLWElvl1& res, cuFHETRGSWNTTlvl1& cs, cuFHETRLWElvl1& c1, cuFHETRLWElvl1& c0, Stream st)
{
    const int d = st.device_id();
    cudaSetDevice(d);
    B2CHECK(b2fhe_defer_fence(d, (b2fhe_stream_t)st.st()));
    B2CHECK(b2fhe_cmux_ntt_batch(d, (b2fhe_stream_t)st.st(), res.trlwedevices[d], cs.trgswdevices[d], c1.trlwedevices[d],
                                 c0.trlwedevices[d], 1, 0));
}
// CMUXNTT (src/cufhe_gates_gpu.cu:69-85): H2D of cs, c1, c0; CMUX; D2H of res
inline void CMUXNTT(cuFHETRLWElvl1& res, cuFHETRGSWNTTlvl1& cs, cuFHETRLWElvl1& c1, cuFHETRLWElvl1& c0, Stream st)
{
    const int d = st.device_id();
    cudaSetDevice(d);
    cudaMemcpyAsync(cs.trgswdevices[d], cs.trgswhost.data(), sizeof(cs.trgswhost), cudaMemcpyHostToDevice, st.st());
    _trlweH2D(c1, st);
    _trlweH2D(c0, st);
    gCMUXNTT(res, cs, c1, c0, st);
    _trlweD2H(res, st);
}

}  // namespace cufhe
