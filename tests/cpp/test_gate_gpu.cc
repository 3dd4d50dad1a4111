// C++ port of the reference's gate test (test/test_gate_gpu.cc:36-91, test/test_util.h:8-96,
// truth tables test/plain.h:10-69) against include/cufhe_b200.hpp -- the SAME user code
// shape: Ctxt<P>, Stream, Nand(out, in0, in1, st) ..., Synchronize(), decrypt == truth.
// Keys/encryption/decryption come from the test oracle (the box has no TFHEpp headers).
// Also exercises the batched calls and checks ciphertext bytes against the oracle.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../include/cufhe_b200.hpp"
#include "../../oracle/cufhe_oracle.h"

using namespace cufhe;
using TFHEpp::lvl0param;
using TFHEpp::lvl1param;

static std::vector<int32_t> sk0(636), sk1(1024);
static int failures = 0;

template <class P> static void enc(TFHEpp::TLWE<P>& c, int bit, uint64_t idx)
{
    if (sizeof(typename P::T) == 2) orc_encrypt_lvl0(7, idx, bit, sk0.data(), (uint16_t*)c.data());
    else orc_encrypt_lvl1(7, idx, bit, sk1.data(), (uint32_t*)c.data());
}
template <class P> static int dec(const TFHEpp::TLWE<P>& c)
{
    return sizeof(typename P::T) == 2 ? orc_decrypt_lvl0((const uint16_t*)c.data(), sk0.data())
                                      : orc_decrypt_lvl1((const uint32_t*)c.data(), sk1.data());
}

template <class P> static void run_level(const TFHEpp::EvalKey& ek, int kNumSMs)
{
    const int level = sizeof(typename P::T) == 2 ? 0 : 1;
    const int kNumTests = 2 * kNumSMs;        // per-gate launches (legacy path)
    std::vector<Stream> st(kNumSMs);
    for (auto& s : st) s.Create();
    std::vector<Ctxt<P>> a(kNumTests), b(kNumTests), c(kNumTests), r(kNumTests);
    std::vector<int> pa(kNumTests), pb(kNumTests), pc(kNumTests);
    for (int op = 0; op < ORC_NUM_OPS; op++) {
        for (int i = 0; i < kNumTests; i++) {
            pa[i] = (i * 7 + op) & 1; pb[i] = (i * 5 + op / 2) & 1; pc[i] = (i * 3 + 1) & 1;
            enc<P>(a[i].tlwehost, pa[i], 1000 * op + i);
            enc<P>(b[i].tlwehost, pb[i], 100000 + 1000 * op + i);
            enc<P>(c[i].tlwehost, pc[i], 200000 + 1000 * op + i);
        }
        for (int i = 0; i < kNumTests; i++) {
            Stream s = st[i % kNumSMs];
            switch (op) {
                case ORC_NAND: Nand<P>(r[i], a[i], b[i], s); break;
                case ORC_OR: Or<P>(r[i], a[i], b[i], s); break;
                case ORC_ORYN: OrYN<P>(r[i], a[i], b[i], s); break;
                case ORC_ORNY: OrNY<P>(r[i], a[i], b[i], s); break;
                case ORC_AND: And<P>(r[i], a[i], b[i], s); break;
                case ORC_ANDYN: AndYN<P>(r[i], a[i], b[i], s); break;
                case ORC_ANDNY: AndNY<P>(r[i], a[i], b[i], s); break;
                case ORC_NOR: Nor<P>(r[i], a[i], b[i], s); break;
                case ORC_XOR: Xor<P>(r[i], a[i], b[i], s); break;
                case ORC_XNOR: Xnor<P>(r[i], a[i], b[i], s); break;
                case ORC_MUX: Mux<P>(r[i], a[i], b[i], c[i], s); break;
                case ORC_NMUX: NMux<P>(r[i], a[i], b[i], c[i], s); break;
                case ORC_NOT: Not<P>(r[i], a[i], s); break;
                case ORC_COPY: Copy<P>(r[i], a[i], s); break;
            }
        }
        Synchronize();
        int bad = 0;
        for (int i = 0; i < kNumTests; i++) {
            if (dec<P>(r[i].tlwehost) != orc_truth(op, pa[i], pb[i], pc[i])) bad++;
            // byte parity with the oracle on a few gates
            if (i < 3) {
                TFHEpp::TLWE<P> want;
                orc_gate(op, level, want.data(), a[i].tlwehost.data(), b[i].tlwehost.data(), c[i].tlwehost.data(),
                         ek.bk.data(), ek.iksk.data());
                if (memcmp(want.data(), r[i].tlwehost.data(), sizeof want)) bad++;
            }
        }
        printf("level %d op %2d : %s (%d gates, per-gate API)\n", level, op, bad ? "FAIL" : "PASS", kNumTests);
        failures += bad;
    }
    // batched API: one launch for 8 x SMs NANDs, chained with a device-resident gNand
    const size_t n = 8 * (size_t)kNumSMs;
    CtxtBatch<P> A(n), B(n), R(n), R2(n);
    std::vector<int> qa(n), qb(n);
    for (size_t i = 0; i < n; i++) {
        qa[i] = (i >> 1) & 1; qb[i] = i & 1;
        enc<P>(A.tlwehost[i], qa[i], 900000 + i);
        enc<P>(B.tlwehost[i], qb[i], 950000 + i);
    }
    Stream s = st[0];
    NandBatch<P>(R, A, B, s);            // H2D + gate + D2H
    gNandBatch<P>(R2, R, B, s);          // device-resident chain: NAND(NAND(a,b), b)
    CtxtCopyD2H<P>(R2, s);
    while (!StreamQuery(s)) {}
    int bad = 0;
    for (size_t i = 0; i < n; i++) {
        const int t1 = 1 - qa[i] * qb[i];
        if (dec<P>(R.tlwehost[i]) != t1) bad++;
        if (dec<P>(R2.tlwehost[i]) != 1 - t1 * qb[i]) bad++;
    }
    printf("level %d batched NAND chain : %s (%zu gates)\n", level, bad ? "FAIL" : "PASS", n);
    failures += bad;
    for (auto& x : st) x.Destroy();
}

// Dependent per-gate calls on device-resident ciphertexts (the reference's g-variants, as a
// circuit built gate by gate would issue them): the deferred queue must respect read-after-
// write, write-after-write and in-place updates; checked by bytes against the oracle, in
// deferred and in immediate mode, then the per-gate API is timed.
template <class P> static void run_chain(const TFHEpp::EvalKey& ek, int kNumSMs, bool deferred)
{
    const int level = sizeof(typename P::T) == 2 ? 0 : 1;
    SetDeferred(deferred);
    const int n = deferred ? 96 : 6;
    std::vector<Stream> st(8);
    for (auto& s : st) s.Create();
    std::vector<Ctxt<P>> a(n), b(n), c(n), x(n), y(n);
    int bad = 0;
    for (int i = 0; i < n; i++) {
        enc<P>(a[i].tlwehost, i & 1, 300000 + i);
        enc<P>(b[i].tlwehost, (i >> 1) & 1, 310000 + i);
        enc<P>(c[i].tlwehost, (i >> 2) & 1, 320000 + i);
    }
    std::vector<TFHEpp::TLWE<P>> wx(n), wy(n), wa(n);
    for (int i = 0; i < n; i++) {
        Stream s = st[i % st.size()];
        CtxtCopyH2D<P>(a[i], s); CtxtCopyH2D<P>(b[i], s); CtxtCopyH2D<P>(c[i], s);
        gNand<P>(x[i], a[i], b[i], s);              // x = NAND(a, b)
        gXor<P>(y[i], x[i], c[i], s);               // y = XOR(x, c)           read after write
        gMux<P>(y[i], y[i], x[i], a[i], s);         // y = y ? x : a           in place, after its writer
        gAnd<P>(a[i], a[i], y[i], s);               // a = AND(a, y)           overwrites an operand of earlier gates
        gNot<P>(x[i], x[i], s);                     // x = NOT(x)              write after read of x by the Mux
        CtxtCopyD2H<P>(x[i], s); CtxtCopyD2H<P>(y[i], s); CtxtCopyD2H<P>(a[i], s);
    }
    for (auto& s : st) while (!StreamQuery(s)) {}
    Synchronize();
    for (int i = 0; i < (deferred ? 8 : n); i++) {
        TFHEpp::TLWE<P> a0, t, u;
        enc<P>(a0, i & 1, 300000 + i);
        orc_gate(ORC_NAND, level, wx[i].data(), a0.data(), b[i].tlwehost.data(), nullptr, ek.bk.data(), ek.iksk.data());
        orc_gate(ORC_XOR, level, t.data(), wx[i].data(), c[i].tlwehost.data(), nullptr, ek.bk.data(), ek.iksk.data());
        orc_gate(ORC_MUX, level, wy[i].data(), t.data(), wx[i].data(), a0.data(), ek.bk.data(), ek.iksk.data());
        orc_gate(ORC_AND, level, wa[i].data(), a0.data(), wy[i].data(), nullptr, ek.bk.data(), ek.iksk.data());
        orc_gate(ORC_NOT, level, u.data(), wx[i].data(), nullptr, nullptr, ek.bk.data(), ek.iksk.data());
        if (memcmp(u.data(), x[i].tlwehost.data(), sizeof u)) bad++;
        if (memcmp(wy[i].data(), y[i].tlwehost.data(), sizeof u)) bad++;
        if (memcmp(wa[i].data(), a[i].tlwehost.data(), sizeof u)) bad++;
    }
    for (int i = 0; i < n; i++) {
        const int pa = i & 1, pb = (i >> 1) & 1, pc = (i >> 2) & 1;
        const int px = 1 - pa * pb, t = px ^ pc, py = t ? px : pa;
        if (dec<P>(x[i].tlwehost) != 1 - px || dec<P>(y[i].tlwehost) != py || dec<P>(a[i].tlwehost) != (pa & py)) bad++;
    }
    printf("level %d dependent per-gate chain (%s) : %s (%d chains x 5 gates)\n", level, deferred ? "deferred" : "immediate",
           bad ? "FAIL" : "PASS", n);
    failures += bad;
    if (deferred) {
        // the reference's benchmark loop (test/test_gate_gpu.cc:68-83): one NAND per call, round-robin streams
        const int mmax = 64 * kNumSMs;
        std::vector<Stream> ss(kNumSMs);
        for (auto& s : ss) s.Create();
        std::vector<Ctxt<P>> p(mmax), q(mmax), r(mmax);
        for (int i = 0; i < mmax; i++) { enc<P>(p[i].tlwehost, i & 1, 400000 + i); enc<P>(q[i].tlwehost, (i >> 1) & 1, 500000 + i); }
        for (int m : {16 * kNumSMs, mmax})
            for (int rep = 0; rep < 2; rep++) {
                const auto t0 = std::chrono::steady_clock::now();
                for (int i = 0; i < m; i++) Nand<P>(r[i], p[i], q[i], ss[i % kNumSMs]);
                Synchronize();
                const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                if (rep) printf("level %d per-gate API, %d NAND calls on %d streams: %.1f ms, %.0f gates/s\n", level, m, kNumSMs, sec * 1e3, m / sec);
            }
        int wrong = 0;
        for (int i = 0; i < mmax; i++) wrong += dec<P>(r[i].tlwehost) != 1 - (i & 1) * ((i >> 1) & 1);
        if (wrong) printf("level %d per-gate API results: FAIL (%d)\n", level, wrong);
        failures += wrong;
        for (auto& s : ss) s.Destroy();
    }
    for (auto& s : st) s.Destroy();
    SetDeferred(true);
}

// TRLWE-level entry points (reference test/test_perf.cc shape: Refresh; plus
// GateBootstrappingTLWE2TRLWElvl01NTT and SampleExtractAndKeySwitch), bytes vs the oracle
static void run_trlwe(const TFHEpp::EvalKey& ek)
{
    Stream s;
    s.Create();
    int bad = 0;
    for (int bit = 0; bit < 2; bit++) {
        Ctxt<lvl0param> in, out;
        cuFHETRLWElvl1 t, t2;
        enc<lvl0param>(in.tlwehost, bit, 777 + bit);
        GateBootstrappingTLWE2TRLWElvl01NTT(t, in, s);
        Synchronize();
        std::vector<uint32_t> want(2 * 1024), want2(2 * 1024);
        orc_blind_rotate_trlwe(want.data(), in.tlwehost.data(), ORC_MU1, ek.bk.data());
        if (memcmp(want.data(), t.trlwehost.data(), 8192)) bad++;
        Refresh(t2, t, s);
        SampleExtractAndKeySwitch(out, t2, s);
        Synchronize();
        orc_refresh(want2.data(), want.data(), ORC_MU1, ek.bk.data(), ek.iksk.data());
        if (memcmp(want2.data(), t2.trlwehost.data(), 8192)) bad++;
        TFHEpp::TLWE<lvl0param> w0;
        orc_sample_extract_keyswitch(w0.data(), want2.data(), ek.iksk.data());
        if (memcmp(w0.data(), out.tlwehost.data(), sizeof w0)) bad++;
        if (dec<lvl0param>(out.tlwehost) != bit) bad++;
    }
    printf("TRLWE-level API (TLWE2TRLWE, Refresh, SampleExtractAndKeySwitch) : %s\n", bad ? "FAIL" : "PASS");
    failures += bad;
    s.Destroy();
}

// operand uploads into buffers with a queued / finished device producer.
//  (1) g* producer (device only, host copy stale), then a COPYING gate: the reference's stream order uploads
//      the HOST bytes, so the host value must win;
//  (2) copying producer, raw cudaDeviceSynchronize (not Synchronize), the caller re-encrypts into the same
//      Ctxt's host copy, then a copying gate: the fresh host value must be used.
template <class P> static void run_host_wins()
{
    const int level = sizeof(typename P::T) == 2 ? 0 : 1;
    Stream s;
    s.Create();
    Ctxt<P> a, b, c, out;
    int bad = 0;
    enc<P>(a.tlwehost, 1, 900001); enc<P>(b.tlwehost, 1, 900002);
    CtxtCopyH2D(a, s); CtxtCopyH2D(b, s);
    enc<P>(c.tlwehost, 1, 900003);                 // host copy of c = 1 and stays 1
    gNand<P>(c, a, b, s);                          // device copy of c = NAND(1, 1) = 0, host untouched
    Nand<P>(out, c, a, s);                         // copying gate: uploads c's HOST bytes (1): NAND(1, 1) = 0
    Synchronize();
    if (dec<P>(out.tlwehost) != 0) bad++;
    Nand<P>(c, a, b, s);                           // copying producer: c = 0 on device and host
    StreamQuery(s);                                // starts the queue
    cudaDeviceSynchronize();                       // raw wait, as test/test_intensive.cc:43
    if (dec<P>(c.tlwehost) != 0) bad++;
    enc<P>(c.tlwehost, 1, 900004);                 // caller re-encrypts into the same Ctxt
    Nand<P>(out, c, a, s);                         // must see the fresh 1: NAND(1, 1) = 0 (stale device 0 would give 1)
    Synchronize();
    if (dec<P>(out.tlwehost) != 0) bad++;
    printf("level %d operand uploads after device producers (host bytes win) : %s\n", level, bad ? "FAIL" : "PASS");
    failures += bad;
    s.Destroy();
}

int main()
{
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { printf("no CUDA device\n"); return 2; }
    const int kNumSMs = prop.multiProcessorCount;
    TFHEpp::EvalKey ek;
    ek.bk.resize(ORC_BK_WORDS);
    ek.iksk.resize(ORC_KSK_WORDS);
    orc_keygen(20261104, sk0.data(), sk1.data(), ek.bk.data(), ek.iksk.data());
    SetGPUNum(1);
    Initialize(ek);
    run_level<lvl1param>(ek, kNumSMs);    // what test/test_gate_gpu.cc runs
    run_level<lvl0param>(ek, kNumSMs);    // what test/test_gate_gpu_multi.cc / test_intensive.cc run
    run_chain<lvl1param>(ek, kNumSMs, true);
    run_chain<lvl0param>(ek, kNumSMs, true);
    run_chain<lvl1param>(ek, kNumSMs, false);
    run_trlwe(ek);
    run_host_wins<lvl0param>();
    run_host_wins<lvl1param>();
    CleanUp();
    printf(failures ? "FAILED (%d)\n" : "ALL PASS\n", failures);
    return failures ? 1 : 0;
}
