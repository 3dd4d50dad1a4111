// CPU race detector for the WHOLE throughput kernel k_blind_rotate_s<4> -- test infrastructure.
//
// tests/emul/emul_sched.cpp checks the CMUX step of one gate; this program runs the real kernel (B2_HOST_TRACE
// build) as one CTA of 512 host threads = FOUR gates side by side, from the table staging and the pre-add /
// modulus switch / test vector through a few CMUX steps to SampleExtract, with three gates present (the fourth
// gate slot of the CTA returns early, as in the last CTA of a batch).  What it adds over the one-gate run: the
// gates' named barriers (ids 1 + gate, 5 + 2 gate + prime) must not interfere, the head and the tail must be
// ordered against the loop, and the gates' shared-memory regions must be disjoint.  Vector-clock race detection
// by tests/emul/trace_rt.h; the lvl1 TLWEs written are compared with the oracle (test vector + CMUX steps +
// SampleExtract(0)).
// usage: emul_brs [drop]   drop = 1..10 removes that synchronisation point of the step in every gate.
#define B2_HOST_TRACE 1
#define B2_BRS_STEPS 3
#include <random>
#include "../../cufhe_b200/csrc/k_blindrotate_s.cuh"
#include "../../oracle/cufhe_oracle.h"
#define TRACE_THREADS 512
#define TRACE_CTA_THREADS 512
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

int main(int argc, char** argv)
{
    trace::drop_site = argc > 1 ? atoi(argv[1]) : 0;
    const int steps = B2_BRS_STEPS, gates = 3;
    static TeamSConst CS;
    make_tables16s(L, CS);
    c_s = CS;
    std::mt19937_64 rng(20261108);
    trace::cta_words = (size_t)brs_smem_bytes(4) / 4;
    trace::smem.assign(trace::cta_words, 0u);
    trace::shadow.resize(trace::smem.size());

    // NAND pre-add of two ciphertext arrays: c = -in0 - in1 + (0, mu0)
    std::vector<uint16_t> in0((size_t)gates * kLvl0Words), in1((size_t)gates * kLvl0Words);
    for (auto& v : in0) v = (uint16_t)rng();
    for (auto& v : in1) v = (uint16_t)rng();
    std::vector<uint32_t> bk_raw((size_t)steps * 8 * kN), bk_dev((size_t)steps * 16 * kN), out((size_t)gates * kLvl1Words, 0xDEADBEEFu);
    for (auto& v : bk_raw) v = (uint32_t)rng();
    for (int i = 0; i < steps; i++)
        for (int row = 0; row < 4; row++)
            for (int q = 0; q < 2; q++)
                for (int o = 0; o < 2; o++)
                    ntt_host(&bk_dev[(((size_t)i * 4 + row) * 2 + q) * 2 * kN + (size_t)o * kN], &bk_raw[((size_t)i * 4 + row) * 2 * kN + (size_t)o * kN], q);

    BrParams P{};
    P.in0 = in0.data(); P.in1 = in1.data(); P.out = out.data(); P.bk = bk_dev.data(); P.count = gates;
    P.ca = -1; P.cb = -1; P.off = (int)kMu0; P.mu = kMu1; P.trlwe = 0; P.first = 0;
    trace::launch(1, [&](int, int) { k_blind_rotate_s<4>(P, reinterpret_cast<const TeamSTables*>(&L)); });

    // the oracle: pre-add, modulus switch, test vector, CMUX steps, SampleExtract(0)
    int bad = 0;
    for (int g = 0; g < gates; g++) {
        static uint32_t ref[2 * kN];
        uint16_t c[kLvl0Words];
        for (int i = 0; i < kLvl0Words; i++) c[i] = (uint16_t)(0u - in0[(size_t)g * kLvl0Words + i] - in1[(size_t)g * kLvl0Words + i]);
        c[kLwe] = (uint16_t)(c[kLwe] + kMu0);
        const uint32_t bbar = modswitch_b(c[kLwe]);
        for (int k = 0; k < kN; k++) { ref[k] = 0u; ref[kN + k] = testvec_b((uint32_t)k, bbar, kMu1); }
        for (int i = 0; i < steps; i++) orc_cmux_step(ref, modswitch_a(c[i]), &bk_raw[(size_t)i * 8 * kN], 0);
        const uint32_t* o = &out[(size_t)g * kLvl1Words];
        for (int m = 0; m < kN; m++) bad += o[m] != (m == 0 ? ref[0] : 0u - ref[kN - m]);
        bad += o[kN] != ref[kN];
    }
    printf("gates %d of 4 per CTA, steps %d, dropped synchronisation point %d: %ld shared-memory accesses traced, races %ld, output words differing from the oracle %d\n",
           gates, steps, trace::drop_site, trace::events, trace::races, bad);
    trace::print_races();
    return 0;
}
