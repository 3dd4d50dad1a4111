// k_blindrotate_cl4.cuh -- lowest-latency blind rotation, second design: one gate per CLUSTER of
// 4 CTAs x 128 threads.  CTA (q, j) = (prime, accumulator polynomial); its two 64-thread teams are
// the two gadget digits d of acc_j, i.e. TRGSW rows r = 2j + d.
//
// k_blind_rotate_cl (8 CTAs x 64 threads) moves 24 KB per step out of the CTAs that run the inverse
// transforms, and distributed shared memory sustains only ~20 B/clk per CTA
// (/opt/skills/guides/B300_MICROARCH.md, "CGA; DSMEM"): a third of its step is transfer + barrier.
// Here a step moves 8 KB per CTA:
//   A. both teams: digits of acc_j (local copy), forward transform, products with their key row for
//      both outputs; the teams add their products through local shared memory, team d finishing
//      output d.  The partial sum of output j stays in registers (team j); the partial sum of output
//      1-j goes to CTA (q, 1-j)                                          [4 KB]  -> cluster barrier
//   B. team j: adds the partial received from CTA (q, 1-j), inverse transform, result to CTA (1-q, j)
//      and to local memory                                               [4 KB]  -> cluster barrier
//   C. all 128 threads: centred CRT of the two primes' results, acc_j += (the CTAs (0, j) and (1, j)
//      keep identical copies of acc_j).
// Same arithmetic (ntt16.cuh), key layout and tables as k_blind_rotate16: bit-identical results.
//
// Replaces (for batches of a few gates) __BlindRotatePreAdd__ + Accumulate
// (reference include/gatebootstrapping_gpu.cuh:115-345), SampleExtract src/bootstrap_gpu.cu:366-381.
#pragma once
#include <cooperative_groups.h>
#include "k_blindrotate16.cuh"

namespace b2 {

constexpr int kCl4Ctas = 4;
constexpr int kCl4Threads = 128;
// tables | acc_j | 2 teams x 2 exchange buffers | 2 team hand-over buffers | received partial |
// my inverse result | the other prime's inverse result | abar
constexpr int kCl4SmemWords = kLane16TableWords + kN + 4 * kXbufWords + 2 * kN + kN + kN + kN + 640 / 2;
// more than half an SM's shared memory, so that no two CTAs of a cluster share an SM
constexpr int kCl4SmemBytes = (kCl4SmemWords * 4 > 120 * 1024) ? kCl4SmemWords * 4 : 120 * 1024;

__global__ void __cluster_dims__(kCl4Ctas, 1, 1) __launch_bounds__(kCl4Threads, 1)
    k_blind_rotate_cl4(const BrParams P, const Lane16Tables* __restrict__ tab16)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    uint32_t* acc = smem + kLane16TableWords;          // acc_j
    uint32_t* xbufs = acc + kN;
    uint32_t* tx = xbufs + 4 * kXbufWords;             // [team][k * 64 + T]: product for the OTHER team's output
    uint32_t* recv = tx + 2 * kN;                      // partial sum of output j from CTA (q, 1-j)
    uint32_t* zloc = recv + kN;                        // inverse result, my prime      [e]
    uint32_t* zrecv = zloc + kN;                       // inverse result, other prime   [e]
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(zrecv + kN);
    stage_tables16(s_tab, tab16);

    const int d = threadIdx.x >> 6, T = threadIdx.x & 63;      // team = gadget digit
    const int rank = (int)cluster.block_rank();
    const int q = rank >> 1, j = rank & 1;
    const int r = 2 * j + d;                                   // TRGSW row
    const int item = blockIdx.x / kCl4Ctas;
    uint32_t* bufA = xbufs + d * 2 * kXbufWords;
    uint32_t* bufB = bufA + kXbufWords;
    const int bar = 1 + d;                                     // named barrier of my team

    // ---- pre-add + modulus switch (gatebootstrapping_gpu.cuh:316-345), redundantly per CTA ----
    BrItem it;
    if (P.items) it = P.items[item];
    else { it.g0 = it.g1 = it.slot = (uint32_t)item; it.ca = (int8_t)P.ca; it.cb = (int8_t)P.cb; it.src1 = 0; it.off = P.off; }
    const uint16_t* a0 = P.in0 + (size_t)it.g0 * kLvl0Words;
    const uint16_t* src1 = it.src1 ? P.in2 : P.in1;
    const uint16_t* a1 = (it.cb != 0 && src1) ? src1 + (size_t)it.g1 * kLvl0Words : nullptr;
    const int ca = it.ca, cb = a1 ? it.cb : 0;
    for (int i = threadIdx.x; i < kLwe; i += kCl4Threads) {
        const int c = ca * (int)a0[i] + (a1 ? cb * (int)a1[i] : 0);
        s_abar[i] = (uint16_t)modswitch_a((uint32_t)c & 0xFFFFu);
    }
    const int cbody = it.off + ca * (int)a0[kLwe] + (a1 ? cb * (int)a1[kLwe] : 0);
    const uint32_t bbar = modswitch_b((uint32_t)cbody & 0xFFFFu);
    for (int c = threadIdx.x; c < kN; c += kCl4Threads) acc[c] = j ? testvec_b((uint32_t)c, bbar, P.mu) : 0u;
    __syncthreads();

    const uint32_t p = q ? Prime<1>::p : Prime<0>::p;
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    const uint32_t dig_off = p - 128u;
    const X16 xa = make_x16(T);
    const uint4* bk4 = reinterpret_cast<const uint4*>(P.bk) + T + (size_t)(r * 2 + q) * 2 * (kN / 4);
    const int dshift = d ? 16 : 24;
    // remote destinations: the partial of output 1-j -> CTA (q, 1-j); my inverse result -> CTA (1-q, j)
    uint32_t* dst_part = cluster.map_shared_rank(recv, q * 2 + (1 - j)) + T;
    uint32_t* dst_z = cluster.map_shared_rank(zrecv, (1 - q) * 2 + j) + T;
    cluster.sync();                                            // every CTA's shared memory is live

    for (int i = 0; i < kLwe; i++) {
        const uint32_t abar = s_abar[i];
        const uint32_t am = abar & (kN - 1);
        const bool flip = (abar >> kNbit) != 0;
        const uint32_t rbase = ((uint32_t)T - abar) & (kN - 1);
        const int kth = ((int)am - T + 63) >> 6;
        const uint4* row = bk4 + (size_t)i * kRows * (4 * (kN / 4));
        uint4 kb[8];                                           // key row first: its latency hides behind the transform
#pragma unroll
        for (int o = 0; o < 2; o++)
#pragma unroll
            for (int v = 0; v < 4; v++) kb[4 * o + v] = __ldg(row + o * (kN / 4) + v * 64);
        // ---- A: digits, forward transform, products -------------------------------------------
        uint32_t x[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            uint32_t rot = acc[(rbase + 64 * k) & (kN - 1)];
            const bool neg = (k < kth) != flip;
            rot = neg ? 0u - rot : rot;
            const uint32_t t = rot - acc[T + 64 * k] + kDecompAdd;
            x[k] = ((t >> dshift) & 0xFFu) + dig_off;
        }
        fwd16(x, bufA, bufB, s_tab, xa, T, q, p, bar, 64);
        uint32_t y[16];                                        // my product for output d (the one my team finishes)
#pragma unroll
        for (int v = 0; v < 4; v++) {
            const uint4 bm = d ? kb[4 + v] : kb[v];            // output d
            const uint4 bo = d ? kb[v] : kb[4 + v];            // output 1 - d: handed to the other team
            y[4 * v + 0] = (uint32_t)umontmul_p(x[4 * v + 0], bm.x, p, pinv) + p;      // (0, 2p)
            y[4 * v + 1] = (uint32_t)umontmul_p(x[4 * v + 1], bm.y, p, pinv) + p;
            y[4 * v + 2] = (uint32_t)umontmul_p(x[4 * v + 2], bm.z, p, pinv) + p;
            y[4 * v + 3] = (uint32_t)umontmul_p(x[4 * v + 3], bm.w, p, pinv) + p;
            uint32_t* t = tx + d * kN + (4 * v) * 64 + T;
            t[0] = (uint32_t)umontmul_p(x[4 * v + 0], bo.x, p, pinv) + p;
            t[64] = (uint32_t)umontmul_p(x[4 * v + 1], bo.y, p, pinv) + p;
            t[128] = (uint32_t)umontmul_p(x[4 * v + 2], bo.z, p, pinv) + p;
            t[192] = (uint32_t)umontmul_p(x[4 * v + 3], bo.w, p, pinv) + p;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; k++) y[k] += tx[(1 - d) * kN + k * 64 + T];            // both rows of acc_j: (0, 4p)
        if (d != j) {                                          // output 1-j is finished by CTA (q, 1-j)
#pragma unroll
            for (int k = 0; k < 16; k++) dst_part[k * 64] = y[k];
        }
        cluster.sync();
        // ---- B: team j sums the two accumulator polynomials' contributions, inverse transform ----
        if (d == j) {
#pragma unroll
            for (int k = 0; k < 16; k++) y[k] += recv[k * 64 + T];                     // (0, 8p)
            inv16(y, bufA, bufB, s_tab, xa, T, q, p, bar, 64);
#pragma unroll
            for (int k = 0; k < 16; k++) {
                zloc[T + 64 * k] = y[k];
                dst_z[64 * k] = y[k];
            }
        }
        cluster.sync();
        // ---- C: CRT of output j, local accumulator update (8 coefficients per thread) ------------
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int e = (int)threadIdx.x + kCl4Threads * k;
            const uint32_t s0 = q ? zrecv[e] : zloc[e];
            const uint32_t s1 = q ? zloc[e] : zrecv[e];
            acc[e] += crt_lift32(s0, s1, c_crt_p0inv.w, c_crt_p0inv.ws);
        }
        __syncthreads();                                       // rotated reads of the next step cross threads
    }

    // ---- tail: CTA (0, 0) holds acc_a, CTA (0, 1) holds acc_b ----
    if (q != 0) return;
    if (P.trlwe) {
        uint32_t* out = P.out + (size_t)it.slot * 2 * kN + j * kN;
        for (int m = threadIdx.x; m < kN; m += kCl4Threads) out[m] = acc[m];
        return;
    }
    uint32_t* out = P.out + (size_t)it.slot * kLvl1Words;      // SampleExtract(0), src/bootstrap_gpu.cu:366-381
    if (j == 0) { for (int m = threadIdx.x; m < kN; m += kCl4Threads) out[m] = m == 0 ? acc[0] : 0u - acc[kN - m]; }
    else if (threadIdx.x == 0) out[kN] = acc[0];
}

}  // namespace b2
