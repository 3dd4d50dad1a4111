// CPU race detector for the stand-alone transform kernels -- test infrastructure.
//
// Runs the REAL kernels of cufhe_b200/csrc/k_ntt_s.cuh (B2_HOST_TRACE build, one persistent CTA, one host thread
// per CUDA thread): k_ntt_fwd_s<0, true> and k_ntt_inv_s<true> with their bulk-copy (TMA) input prefetch -- a
// staging buffer that is re-armed and overwritten while other threads may still read it, an mbarrier with a
// transaction count --, k_polymul_s<2, 5> / <3, 3> and the stand-alone CMUX k_cmux_s<4> (two of the four item slots used).  tests/emul/trace_rt.h reports every pair of unordered
// conflicting shared-memory accesses; the results are checked as well: INTT(NTT(x)) == x and the products
// against the oracle's schoolbook product.
// usage: emul_nttk [polys] [drop] [deadlock seconds]   drop = 11 / 12 / 13 removes the barrier that protects the staging buffer of the
// forward / inverse / product kernel from being refilled while it is read; the detector must then report (races, and the
// deadlock that follows when the mbarrier's phase runs ahead of a waiting thread).
#define B2_HOST_TRACE 1
#include <random>
#include "../../cufhe_b200/csrc/k_cmux.cuh"
#include "../../oracle/cufhe_oracle.h"
#define TRACE_THREADS 640
#define TRACE_CTA_THREADS 640
#include "trace_rt.h"

using namespace b2;
static TeamSTables3 L;

static void reset(size_t bytes)
{
    trace::cta_words = bytes / 4;
    trace::smem.assign(trace::cta_words, 0u);
    trace::shadow.assign(trace::smem.size(), trace::Shadow{});
    trace::bars.clear();
    trace::mbars.clear();
    for (auto& v : trace::vc) v = trace::VC{};
}

int main(int argc, char** argv)
{
    const int n = argc > 1 ? atoi(argv[1]) : 13;          // ragged: the four polynomial slots of the CTA take 4, 3, 3, 3
    trace::drop_site = argc > 2 ? atoi(argv[2]) : 0;
    if (argc > 3) trace::deadlock_s = atoi(argv[3]);
    static TeamSConst CS;
    make_tables16s(L, CS);
    c_s = CS;
    std::mt19937_64 rng(20261107);
    std::vector<uint32_t> a((size_t)n * kN), b((size_t)n * kN), res((size_t)n * 2 * kN), back((size_t)n * kN), prod((size_t)n * kN);
    for (auto& v : a) v = (uint32_t)rng();
    for (auto& v : b) v = (uint32_t)(rng() % (1u << 18));

    long races = 0, events = 0;
    // forward, inputs prefetched by bulk copies
    reset((size_t)ntt_fwd_pf_smem_bytes<kNttTeams>());
    trace::launch(1, [&](int, int) { k_ntt_fwd_s<0, true>(a.data(), res.data(), n, L.pr); }, 512);
    printf("k_ntt_fwd_s<0, true>: %ld accesses, races %ld\n", trace::events, trace::races);
    races += trace::races; events += trace::events; trace::races = trace::events = 0;
    // inverse, inputs prefetched by bulk copies
    reset((size_t)ntt_inv_pf_smem_bytes<kNttTeams>());
    trace::launch(1, [&](int, int) { k_ntt_inv_s<true>(res.data(), back.data(), n, L.pr); }, 512);
    printf("k_ntt_inv_s<true>:    %ld accesses, races %ld\n", trace::events, trace::races);
    races += trace::races; events += trace::events; trace::races = trace::events = 0;
    int bad_rt = 0;
    for (size_t i = 0; i < a.size(); i++) bad_rt += back[i] != a[i];
    int bad_range = 0;
    for (int p = 0; p < n; p++)
        for (int q = 0; q < 2; q++)
            for (int i = 0; i < kN; i++) bad_range += res[((size_t)p * 2 + q) * kN + i] >= kP[q];

    // products: two primes (b small enough for P ~ 2^53: 8-bit), three primes (b < 2^18, the reference test's domain)
    int bad_prod = 0;
    {
        std::vector<uint32_t> b8((size_t)n * kN);
        for (auto& v : b8) v = (uint32_t)(int32_t)((int)(rng() % 256) - 128);
        reset((size_t)polymul_pf_smem_bytes<2, 5>());
        trace::launch(1, [&](int, int) { k_polymul_s<2, 5, true>(a.data(), b8.data(), prod.data(), n, L.pr); }, 640);
        printf("k_polymul_s<2, 5, 1>: %ld accesses, races %ld\n", trace::events, trace::races);
        races += trace::races; events += trace::events; trace::races = trace::events = 0;
        static uint32_t ref[kN];
        for (int p = 0; p < n; p++) {
            orc_polymul_schoolbook(ref, &a[(size_t)p * kN], &b8[(size_t)p * kN]);
            bad_prod += memcmp(ref, &prod[(size_t)p * kN], sizeof ref) != 0;
        }
        reset((size_t)polymul_pf_smem_bytes<3, 3>());
        trace::launch(1, [&](int, int) { k_polymul_s<3, 3, true>(a.data(), b.data(), prod.data(), n, L.pr); }, 576);
        printf("k_polymul_s<3, 3, 1>: %ld accesses, races %ld\n", trace::events, trace::races);
        races += trace::races; events += trace::events;
        for (int p = 0; p < n; p++) {
            orc_polymul_schoolbook(ref, &a[(size_t)p * kN], &b[(size_t)p * kN]);
            bad_prod += memcmp(ref, &prod[(size_t)p * kN], sizeof ref) != 0;
        }
    }
    // stand-alone CMUX: out = c0 + (c1 - c0) [.] TRGSW, two items in a four-item CTA, one TRGSW per item
    int bad_cmux = 0;
    {
        const int items = 2;
        std::vector<uint32_t> tr_raw((size_t)items * 8 * kN), tr_dev((size_t)items * 16 * kN), c1((size_t)items * 2 * kN), c0((size_t)items * 2 * kN), co((size_t)items * 2 * kN);
        for (auto& v : tr_raw) v = (uint32_t)rng();
        for (auto& v : c1) v = (uint32_t)rng();
        for (auto& v : c0) v = (uint32_t)rng();
        auto ntt_host = [&](uint32_t* dst, const uint32_t* src, int q) {
            const uint32_t p = kP[q], psi = h_psi(p);
            std::vector<uint32_t> w(kN);
            for (int i = 0; i < kN; i++) w[i] = src[i] % p;
            for (int s = 0, len = kN / 2; s < kNbit; s++, len >>= 1)
                for (int g = 0; g < (1 << s); g++) {
                    const uint32_t tw = h_powmod(psi, brv10((1u << s) + g), p);
                    for (int j = 0; j < len; j++) {
                        const int u = g * 2 * len + j, v = u + len;
                        const uint32_t t = h_mulmod(w[v], tw, p);
                        w[v] = (w[u] + p - t) % p;
                        w[u] = (w[u] + t) % p;
                    }
                }
            const uint32_t r32 = (uint32_t)((1ull << 32) % p);
            for (int T = 0; T < 64; T++)
                for (int k = 0; k < 16; k++) dst[bk16_slot(T, k)] = h_mulmod(w[eC(T, k)], r32, p);
        };
        for (int it = 0; it < items; it++)
            for (int row = 0; row < 4; row++)
                for (int q = 0; q < 2; q++)
                    for (int o = 0; o < 2; o++)
                        ntt_host(&tr_dev[(((size_t)it * 4 + row) * 2 + q) * 2 * kN + (size_t)o * kN], &tr_raw[((size_t)it * 4 + row) * 2 * kN + (size_t)o * kN], q);
        CmuxParams CP{};
        CP.cs = tr_dev.data(); CP.c1 = c1.data(); CP.c0 = c0.data(); CP.out = co.data(); CP.count = items; CP.cs_per_item = 1;
        trace::races = trace::events = 0;
        reset((size_t)cmux_smem_bytes(4));
        trace::launch(1, [&](int, int) { k_cmux_s<4>(CP, L.pr); }, 512);
        printf("k_cmux_s<4>:          %ld accesses, races %ld\n", trace::events, trace::races);
        races += trace::races; events += trace::events;
        static uint32_t ref[2 * kN];
        for (int it = 0; it < items; it++) {
            orc_cmux(ref, &tr_raw[(size_t)it * 8 * kN], &c1[(size_t)it * 2 * kN], &c0[(size_t)it * 2 * kN], 0);
            bad_cmux += memcmp(ref, &co[(size_t)it * 2 * kN], sizeof ref) != 0;
        }
    }
    bad_prod += bad_cmux;
    printf("polynomials %d: %ld shared-memory accesses traced, races %ld, round-trip words wrong %d, residues out of range %d, products / CMUX results wrong %d\n",
           n, events, races, bad_rt, bad_range, bad_prod);
    trace::print_races();
    return 0;
}
