// CPU race detector for the CMUX step schedule of k_blind_rotate_s -- test infrastructure.
//
// compute-sanitizer (racecheck / synccheck) is closed on the GPU pool, and timing perturbation on the GPU
// (tools/fuzz_mutants.cu) cannot expose a missing __syncwarp().  This program runs the REAL source of the step
// (brs_gate_loop and everything it calls in cufhe_b200/csrc/k_blindrotate_s.cuh + ntt16s.cuh, compiled for the
// host with B2_HOST_TRACE) with one host thread per CUDA thread of a gate (128), every lane an independent thread:
//   * shared memory is an emulated array; every ld.shared / st.shared goes through trace_ld / trace_st,
//   * bar.sync id, n and __syncwarp() are real blocking barriers between the host threads,
//   * happens-before is tracked with vector clocks (a barrier joins the clocks of the threads it releases); two
//     accesses to the same 32-bit word from different threads, at least one a store, that are not ordered by
//     barriers are reported as a race -- whatever order the host scheduler happened to run them in,
//   * the accumulator after the steps is compared with the oracle's CMUX steps (the real kernel arithmetic
//     runs here, on real tables and a real NTT-domain key).
// usage: emul_sched [steps] [drop_site]   drop_site = 1..10 removes that synchronisation point of the step
// (the numbering of tbar_sync<ABL, SITE> / warp_sync<ABL, SITE>); the detector must then report races.
#define B2_HOST_TRACE 1
#include <random>
#include "../../cufhe_b200/csrc/k_blindrotate_s.cuh"
#include "../../oracle/cufhe_oracle.h"
#define TRACE_THREADS 128                       // one gate: two teams of 64
#define TRACE_CTA_THREADS 128
#include "trace_rt.h"

using namespace b2;

// ---- key conversion on the host: raw TRGSW polynomial -> NTT domain, Montgomery form, bk16_slot order ----
static TeamSTables3 L;
static void ntt_host(uint32_t* dst, const uint32_t* src, int q)
{   // plain O(N^2)-free reference transform: evaluate at the positions the team schedule produces, via the schedule's own
    // index convention (merged-psi CT, natural -> bit-reversed): position e holds the value at psi^(2 brv(e) + 1)
    const uint32_t p = kP[q];
    const uint32_t psi = h_psi(p);
    static uint32_t a[kN];
    for (int i = 0; i < kN; i++) a[i] = src[i] % p;
    // iterative CT with the same twiddle rule as ntt16s.cuh (stage s, group i: psi^brv(2^s + i))
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
        for (int k = 0; k < 16; k++) dst[bk16_slot(T, k)] = h_mulmod(a[eC(T, k)], r32, p);      // Montgomery form
}

int main(int argc, char** argv)
{
    const int steps = argc > 1 ? atoi(argv[1]) : 2;
    trace::drop_site = argc > 2 ? atoi(argv[2]) : 0;
    static TeamSConst CS;
    make_tables16s(L, CS);
    c_s = CS;
    std::mt19937_64 rng(20261105);

    // shared memory of a one-gate CTA, laid out as the kernel lays it out
    trace::cta_words = (size_t)brs_smem_bytes(1) / 4;
    trace::smem.assign(trace::cta_words, 0u);
    trace::shadow.resize(trace::smem.size());
    memcpy(trace::smem.data(), &L, sizeof(TeamSTables));                    // the first two primes
    uint32_t* gbase = trace::smem.data() + kTeamSTableWords;
    uint32_t* acc = gbase;                                                  // [-acc_0 | +acc_0 | -acc_1 | +acc_1]
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(gbase + kBrsAccWords + 4 * kXsWords);
    static uint32_t ref[2 * kN];
    for (int j = 0; j < 2; j++)
        for (int c = 0; c < kN; c++) {
            const uint32_t v = (uint32_t)rng();
            ref[j * kN + c] = v;
            acc[2 * j * kN + c] = 0u - v;
            acc[(2 * j + 1) * kN + c] = v;
        }
    // key rows of the steps: raw [4 rows][2 polys][1024] -> device layout [row][prime][out][1024]
    std::vector<uint32_t> bk_raw((size_t)steps * 8 * kN), bk_dev((size_t)steps * 16 * kN);
    for (auto& v : bk_raw) v = (uint32_t)rng();
    for (int i = 0; i < steps; i++) {
        s_abar[i] = (uint16_t)(i == 0 ? 1500 : rng() % 2048);               // one step with a >= 1024 (the flip branch)
        for (int row = 0; row < 4; row++)
            for (int q = 0; q < 2; q++)
                for (int o = 0; o < 2; o++)
                    ntt_host(&bk_dev[(((size_t)i * 4 + row) * 2 + q) * 2 * kN + (size_t)o * kN], &bk_raw[((size_t)i * 4 + row) * 2 * kN + (size_t)o * kN], q);
        orc_cmux_step(ref, s_abar[i], &bk_raw[(size_t)i * 8 * kN], 0);
    }

    const TeamSTables* tab = reinterpret_cast<const TeamSTables*>(trace::smem.data());
    trace::launch(1, [&](int, int t) {
        const int q = t >> 6, T = t & 63;
        uint32_t* xs = gbase + kBrsAccWords + q * 2 * kXsWords;
        uint32_t* xs_other = gbase + kBrsAccWords + (q ^ 1) * 2 * kXsWords;
        const SAddr a = brs_addr(tab, xs, xs_other, q, T);
        const int gbar = 1, tbar = 2 + q;
        brs_gate_loop<false, 0, 0>(bk_dev.data(), smem_u32(acc), a, s_abar, T, q, tbar, gbar, 0u, steps);
    });

    int bad = 0;
    for (int j = 0; j < 2; j++)
        for (int c = 0; c < kN; c++) {
            bad += acc[(2 * j + 1) * kN + c] != ref[j * kN + c];
            bad += acc[2 * j * kN + c] != 0u - ref[j * kN + c];
        }
    printf("steps %d, dropped synchronisation point %d: %ld shared-memory accesses traced, races %ld, accumulator words differing from the oracle %d\n",
           steps, trace::drop_site, trace::events, trace::races, bad);
    trace::print_races();
    return 0;
}
