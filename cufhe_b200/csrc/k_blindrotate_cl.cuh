// k_blindrotate_cl.cuh -- LOWEST-LATENCY blind rotation: one gate per CLUSTER of 8 CTAs.
//
// k_blind_rotate_lat keeps a gate on one SM, where its 16 warps still share that SM's
// integer-multiply pipes (ncu/probe: ~50 % of a step is IMAD-pipe time).  Here the eight
// 64-thread teams (prime q, TRGSW row r) of a gate are eight CTAs of a thread-block cluster,
// each alone on its own SM, and talk through distributed shared memory:
//   1. every CTA: digits of its row from its LOCAL copy of the accumulator, one forward
//      transform, products with its key row, stored straight into the shared memory of the
//      CTA that owns (prime q, output o)                                  -> cluster barrier
//   2. CTAs (q, o = r < 2): sum the four row products, one inverse transform, result
//      broadcast to the four CTAs whose digits come from output o       -> cluster barrier
//   3. every CTA: centred CRT of the ONE accumulator polynomial it decomposes (rows 2j, 2j+1
//      need acc_j only) and update of its local copy (redundant across 4 CTAs, deterministic
//      -- cheaper than a third barrier).
// Two cluster barriers per CMUX step; same arithmetic (ntt16.cuh), key layout and tables as
// k_blind_rotate16: bit-identical results.
//
// Replaces (for batches of a few gates) __BlindRotatePreAdd__ + Accumulate
// (reference include/gatebootstrapping_gpu.cuh:115-345), SampleExtract src/bootstrap_gpu.cu:366-381.
#pragma once
#include <cooperative_groups.h>
#include "k_blindrotate16.cuh"

namespace b2 {

constexpr int kClCtas = 8;
// tables | acc_j | 2 exchange buffers | 4 row products of my output | 2 primes' inverse results of output j | abar
constexpr int kClSmemWords = kLane16TableWords + kN + 2 * kXbufWords + 4 * kN + 2 * kN + 640 / 2;
// ask for more than half an SM's shared memory so that no two CTAs of a cluster share an SM
constexpr int kClSmemBytes = (kClSmemWords * 4 > 120 * 1024) ? kClSmemWords * 4 : 120 * 1024;

__global__ void __cluster_dims__(kClCtas, 1, 1) __launch_bounds__(64, 1)
    k_blind_rotate_cl(const BrParams P, const Lane16Tables* __restrict__ tab16)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    uint32_t* acc = smem + kLane16TableWords;
    uint32_t* bufA = acc + kN;                         // acc = the ONE polynomial acc_j this CTA decomposes
    uint32_t* bufB = bufA + kXbufWords;
    uint32_t* recv = bufB + kXbufWords;               // [row][k][T]: products for (my prime, output r)
    uint32_t* yall = recv + 4 * kN;                   // [prime][e]: inverse results of output j of this step
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(yall + 2 * kN);
    stage_tables16(s_tab, tab16);

    const int T = threadIdx.x;
    const int rank = (int)cluster.block_rank();
    const int q = rank >> 2, r = rank & 3;            // prime, TRGSW row (r = 2 j + d)
    const int item = blockIdx.x / kClCtas;

    // ---- pre-add + modulus switch (gatebootstrapping_gpu.cuh:316-345), redundantly per CTA ----
    BrItem it;
    if (P.items) it = P.items[item];
    else { it.g0 = it.g1 = it.slot = (uint32_t)item; it.ca = (int8_t)P.ca; it.cb = (int8_t)P.cb; it.src1 = 0; it.off = P.off; }
    const uint16_t* a0 = P.in0 + (size_t)it.g0 * kLvl0Words;
    const uint16_t* src1 = it.src1 ? P.in2 : P.in1;
    const uint16_t* a1 = (it.cb != 0 && src1) ? src1 + (size_t)it.g1 * kLvl0Words : nullptr;
    const int ca = it.ca, cb = a1 ? it.cb : 0;
    for (int i = T; i < kLwe; i += 64) {
        const int c = ca * (int)a0[i] + (a1 ? cb * (int)a1[i] : 0);
        s_abar[i] = (uint16_t)modswitch_a((uint32_t)c & 0xFFFFu);
    }
    const int cbody = it.off + ca * (int)a0[kLwe] + (a1 ? cb * (int)a1[kLwe] : 0);
    const uint32_t bbar = modswitch_b((uint32_t)cbody & 0xFFFFu);
    const int j = rank >> 1 & 1;                       // (r >> 1): which accumulator polynomial I hold
    for (int c = T; c < kN; c += 64) acc[c] = j ? testvec_b((uint32_t)c, bbar, P.mu) : 0u;
    __syncthreads();

    const uint32_t p = q ? Prime<1>::p : Prime<0>::p;
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    const uint32_t dig_off = p - 128u;
    const X16 xa = make_x16(T);
    const uint4* bk4 = reinterpret_cast<const uint4*>(P.bk) + T + (size_t)(r * 2 + q) * 2 * (kN / 4);
    const uint32_t* aj = acc;
    const int dshift = (r & 1) ? 16 : 24;
    // where my products go: CTA (q, o) for o = 0, 1, slot [row r]
    uint32_t* dst_part[2];
#pragma unroll
    for (int o = 0; o < 2; o++) dst_part[o] = cluster.map_shared_rank(recv, q * 4 + o) + r * kN + T;
    // where an inverse result of output o = r goes: yall[q] of the four CTAs (q', 2o), (q', 2o+1)
    uint32_t* dst_y[4];
#pragma unroll
    for (int c = 0; c < 4; c++) dst_y[c] = cluster.map_shared_rank(yall, (c >> 1) * 4 + 2 * (r & 1) + (c & 1)) + q * kN + T;
    cluster.sync();                                   // every CTA's shared memory is live

    for (int i = 0; i < kLwe; i++) {
        const uint32_t abar = s_abar[i];
        const uint32_t am = abar & (kN - 1);
        const bool flip = (abar >> kNbit) != 0;
        const uint32_t rbase = ((uint32_t)T - abar) & (kN - 1);
        const int kth = ((int)am - T + 63) >> 6;
        const uint4* row = bk4 + (size_t)i * kRows * (4 * (kN / 4));
        uint4 kb[8];                                  // key row first: its L2 latency hides behind the transform
#pragma unroll
        for (int o = 0; o < 2; o++)
#pragma unroll
            for (int v = 0; v < 4; v++) kb[4 * o + v] = __ldg(row + o * (kN / 4) + v * 64);
        uint32_t x[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            uint32_t rot = aj[(rbase + 64 * k) & (kN - 1)];
            const bool neg = (k < kth) != flip;
            rot = neg ? 0u - rot : rot;
            const uint32_t t = rot - aj[T + 64 * k] + kDecompAdd;
            x[k] = ((t >> dshift) & 0xFFu) + dig_off;
        }
        fwd16(x, bufA, bufB, s_tab, xa, T, q, p, 1, 64);
#pragma unroll
        for (int o = 0; o < 2; o++)
#pragma unroll
            for (int v = 0; v < 4; v++) {
                const uint4 b = kb[4 * o + v];
                uint32_t* d = dst_part[o] + (4 * v) * 64;
                d[0] = (uint32_t)umontmul_p(x[4 * v + 0], b.x, p, pinv) + p;      // (0, 2p)
                d[64] = (uint32_t)umontmul_p(x[4 * v + 1], b.y, p, pinv) + p;
                d[128] = (uint32_t)umontmul_p(x[4 * v + 2], b.z, p, pinv) + p;
                d[192] = (uint32_t)umontmul_p(x[4 * v + 3], b.w, p, pinv) + p;
            }
        cluster.sync();
        if (r < 2) {
            uint32_t y[16];
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const uint32_t* s = recv + k * 64 + T;
                y[k] = s[0] + s[kN] + s[2 * kN] + s[3 * kN];                      // (0, 8p)
            }
            inv16(y, bufA, bufB, s_tab, xa, T, q, p, 1, 64);
#pragma unroll
            for (int c = 0; c < 4; c++)
#pragma unroll
                for (int k = 0; k < 16; k++) dst_y[c][64 * k] = y[k];
        }
        cluster.sync();
        // every CTA: CRT of output j, local accumulator update (thread owns e = T + 64 k)
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const int e = T + 64 * k;
            acc[e] += crt_lift32(yall[e], yall[kN + e], c_crt_p0inv.w, c_crt_p0inv.ws);
        }
        named_bar_sync(1, 64);                        // rotated reads of the next step cross threads
    }

    // ---- tail: CTA 0 holds acc_a, CTA 2 holds acc_b ----
    if (rank != 0 && rank != 2) return;
    if (P.trlwe) {
        uint32_t* out = P.out + (size_t)it.slot * 2 * kN + j * kN;
        for (int m = T; m < kN; m += 64) out[m] = acc[m];
        return;
    }
    uint32_t* out = P.out + (size_t)it.slot * kLvl1Words;   // SampleExtract(0), src/bootstrap_gpu.cu:366-381
    if (j == 0) { for (int m = T; m < kN; m += 64) out[m] = m == 0 ? acc[0] : 0u - acc[kN - m]; }
    else if (T == 0) out[kN] = acc[0];
}

}  // namespace b2
