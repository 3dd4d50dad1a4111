// CPU race detector for the cluster kernel k_blind_rotate_cl4s -- test infrastructure.
//
// Runs the REAL kernel source (cufhe_b200/csrc/k_blindrotate_cl4s.cuh, B2_HOST_TRACE build: 4 CTAs x 128
// threads = 512 host threads, every lane independent) over an emulated shared memory of the four CTAs of a
// cluster.  mapa, st.async ... mbarrier::complete_tx::bytes, mbarrier arrive.expect_tx / try_wait.parity,
// __syncthreads, the teams' named barriers, __syncwarp and cluster.sync are modelled by tests/emul/trace_rt.h;
// the detector reports every pair of unordered conflicting accesses (local or through distributed shared memory)
// and every mbarrier arrival on a phase that has no arrival left.  The TRLWE the kernel writes after a few CMUX
// steps is compared with the oracle's.
// usage: emul_cl4s [drop]   drop = 1..5 removes that in-loop synchronisation point (the MUT template argument:
// 1, 3, 5 = the three __syncthreads, 2, 4 = the two mbarrier waits); the detector must then report.
#define B2_HOST_TRACE 1
#define B2_CL4_STEPS 5
#include <random>
#include "../../cufhe_b200/csrc/k_blindrotate_cl4s.cuh"
#include "../../oracle/cufhe_oracle.h"
#define TRACE_THREADS 512
#define TRACE_CTA_THREADS 128
#include "trace_rt.h"

using namespace b2;

static TeamSTables3 L;
static void ntt_host(uint32_t* dst, const uint32_t* src, int q)     // raw polynomial -> NTT domain, Montgomery form, bk16_slot order
{
    const uint32_t p = kP[q];
    const uint32_t psi = h_psi(p);
    static uint32_t a[kN];
    for (int i = 0; i < kN; i++) a[i] = src[i] % p;
    for (int s = 0, len = kN / 2; s < kNbit; s++, len >>= 1)
        for (int g = 0; g < (1 << s); g++) {
            const uint32_t w = h_powmod(psi, brv10((1u << s) + g), p);
            for (int j = 0; j < len; j++) {
                const int u = g * 2 * len + j, v = u + len;
                const uint32_t t = h_mulmod(a[v], w, p);
                a[v] = (a[u] + p - t) % p;
                a[u] = (a[u] + t) % p;
            }
        }
    const uint32_t r32 = (uint32_t)((1ull << 32) % p);
    for (int T = 0; T < 64; T++)
        for (int k = 0; k < 16; k++) dst[bk16_slot(T, k)] = h_mulmod(a[eC(T, k)], r32, p);
}

template <int MUT> static void run(const BrParams& P) { trace::launch(kCl4Ctas, [&](int, int) { k_blind_rotate_cl4s<false, MUT>(P, L.pr); }); }

int main(int argc, char** argv)
{
    const int drop = argc > 1 ? atoi(argv[1]) : 0;
    const int steps = kCl4Steps;
    static TeamSConst CS;
    make_tables16s(L, CS);
    c_s = CS;
    std::mt19937_64 rng(20261106);

    trace::cta_words = (size_t)kCl4SmemBytes / 4;
    trace::smem.assign(trace::cta_words * kCl4Ctas, 0u);
    trace::shadow.resize(trace::smem.size());

    // one lvl0 ciphertext; the first mask coefficient is chosen so that step 0 rotates by a >= 1024
    static uint16_t ct[kLvl0Words];
    for (auto& v : ct) v = (uint16_t)rng();
    ct[0] = (uint16_t)(1500u << 5);
    std::vector<uint32_t> bk_raw((size_t)(steps + 1) * 8 * kN), bk_dev((size_t)(steps + 1) * 16 * kN);
    for (auto& v : bk_raw) v = (uint32_t)rng();
    for (int i = 0; i <= steps; i++)                                  // the kernel fetches one step ahead
        for (int row = 0; row < 4; row++)
            for (int q = 0; q < 2; q++)
                for (int o = 0; o < 2; o++)
                    ntt_host(&bk_dev[(((size_t)i * 4 + row) * 2 + q) * 2 * kN + (size_t)o * kN], &bk_raw[((size_t)i * 4 + row) * 2 * kN + (size_t)o * kN], q);

    // the oracle: test vector, then the CMUX steps
    static uint32_t ref[2 * kN], out[2 * kN];
    const uint32_t bbar = modswitch_b(ct[kLwe]);
    for (int c = 0; c < kN; c++) { ref[c] = 0u; ref[kN + c] = testvec_b((uint32_t)c, bbar, kMu1); }
    for (int i = 0; i < steps; i++) orc_cmux_step(ref, modswitch_a(ct[i]), &bk_raw[(size_t)i * 8 * kN], 0);

    BrParams P{};
    P.in0 = ct; P.out = out; P.bk = bk_dev.data(); P.count = 1; P.ca = 1; P.cb = 0; P.off = 0; P.mu = kMu1; P.trlwe = 1;
    switch (drop) {
        case 1: run<1>(P); break;
        case 2: run<2>(P); break;
        case 3: run<3>(P); break;
        case 4: run<4>(P); break;
        case 5: run<5>(P); break;
        default: run<0>(P);
    }
    int bad = 0;
    for (int i = 0; i < 2 * kN; i++) bad += out[i] != ref[i];
    printf("steps %d, dropped synchronisation point %d: %ld shared-memory accesses traced, races %ld, TRLWE words differing from the oracle %d\n",
           steps, drop, trace::events, trace::races, bad);
    trace::print_races();
    return 0;
}
