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
// ASYNC = true: the two hand-overs use st.async stores that complete on an mbarrier of the receiving
// CTA (transaction bytes) instead of "remote stores, cluster-scope fence, cluster barrier": ncu shows
// the barrier version spending 18 % of its warp time in the fence and 17 % at barriers.  Receive
// buffers are double-buffered by step parity: a sender is never two steps ahead of its receiver
// (its own progress needs data the receiver only sends after consuming the previous buffer).
//
// Replaces (for batches of a few gates) __BlindRotatePreAdd__ + Accumulate
// (reference include/gatebootstrapping_gpu.cuh:115-345), SampleExtract src/bootstrap_gpu.cu:366-381.
#pragma once
#include <cooperative_groups.h>
#include "team16_dev.cuh"

namespace b2 {

__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, int rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// default (.acquire.cta) wait, as for TMA-written shared memory: the st.async data is in shared memory
// when the transaction count completes; an .acquire.cluster wait would add an L1 invalidate (CCTL.IVALL)
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MBAR_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra MBAR_DONE_%=;\n"
        "bra MBAR_WAIT_%=;\n"
        "MBAR_DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
// 32-bit store into another CTA's shared memory that signals `bytes written` on that CTA's mbarrier
__device__ __forceinline__ void st_async_u32(uint32_t remote_addr, uint32_t v, uint32_t remote_bar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.u32 [%0], %1, [%2];" ::"r"(remote_addr), "r"(v),
                 "r"(remote_bar) : "memory");
}

constexpr int kCl4Ctas = 4;
constexpr int kCl4Threads = 128;
// tables | acc_j | 2 teams x 2 exchange buffers | 2 team hand-over buffers | received partial |
// my inverse result | the other prime's inverse result | abar
// (ASYNC: received partial and the other prime's result are double-buffered, + 4 mbarriers)
constexpr int kCl4SmemWords = kLane16TableWords + kN + 4 * kXbufWords + 2 * kN + 2 * kN + kN + 2 * kN + 640 / 2 + 8;
// 110 KB: two CTAs (of different gates, or of one gate when few are in flight) fit one SM
constexpr int kCl4SmemBytes = kCl4SmemWords * 4;

template <bool ASYNC>
__global__ void __cluster_dims__(kCl4Ctas, 1, 1) __launch_bounds__(kCl4Threads, 2)
    k_blind_rotate_cl4(const BrParams P, const Lane16Tables* __restrict__ tab16)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    uint32_t* acc = smem + kLane16TableWords;          // acc_j
    uint32_t* xbufs = acc + kN;
    uint32_t* tx = xbufs + 4 * kXbufWords;             // [team][k * 64 + T]: product for the OTHER team's output
    uint32_t* recv = tx + 2 * kN;                      // [parity] partial sum of output j from CTA (q, 1-j)
    uint32_t* zloc = recv + 2 * kN;                    // inverse result, my prime      [e]
    uint32_t* zrecv = zloc + kN;                       // [parity] inverse result, other prime   [e]
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(zrecv + 2 * kN);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(zrecv + 2 * kN + 640 / 2);      // [0..1] recv, [2..3] zrecv
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
    const uint32_t ra_part = mapa_u32(smem_u32(recv + T), q * 2 + (1 - j));
    const uint32_t ra_z = mapa_u32(smem_u32(zrecv + T), (1 - q) * 2 + j);
    const uint32_t rb_part = mapa_u32(smem_u32(mbar), q * 2 + (1 - j));           // + 8 * parity
    const uint32_t rb_z = mapa_u32(smem_u32(mbar + 2), (1 - q) * 2 + j);
    if (ASYNC && threadIdx.x == 0) {
        for (int b = 0; b < 4; b++) mbar_init(smem_u32(mbar + b), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    cluster.sync();                                            // every CTA's shared memory (and mbarriers) are live

    uint4 kb[8];                                               // my key row, fetched ONE STEP AHEAD (L2 latency ~ 1/5 of a step)
#pragma unroll
    for (int o = 0; o < 2; o++)
#pragma unroll
        for (int v = 0; v < 4; v++) kb[4 * o + v] = __ldg(bk4 + o * (kN / 4) + v * 64);
    for (int i = 0; i < kLwe; i++) {
        const uint32_t abar = s_abar[i];
        const uint32_t am = abar & (kN - 1);
        const bool flip = (abar >> kNbit) != 0;
        const uint32_t rbase = ((uint32_t)T - abar) & (kN - 1);
        const int kth = ((int)am - T + 63) >> 6;
        const int par = i & 1;
        const uint32_t phase = (uint32_t)(i >> 1) & 1u;
        if (ASYNC && threadIdx.x == 0) {                       // arm this step's two receive barriers: 4 KB each
            mbar_arrive_expect_tx(smem_u32(mbar + par), kN * 4);
            mbar_arrive_expect_tx(smem_u32(mbar + 2 + par), kN * 4);
        }
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
        {   // the products above were the last use of this step's key row: fetch the next one now
            const uint4* row = bk4 + (size_t)(i + 1 < kLwe ? i + 1 : i) * kRows * (4 * (kN / 4));
#pragma unroll
            for (int o = 0; o < 2; o++)
#pragma unroll
                for (int v = 0; v < 4; v++) kb[4 * o + v] = __ldg(row + o * (kN / 4) + v * 64);
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; k++) y[k] += tx[(1 - d) * kN + k * 64 + T];            // both rows of acc_j: (0, 4p)
        if (d != j) {                                          // output 1-j is finished by CTA (q, 1-j)
            if (ASYNC) {
#pragma unroll
                for (int k = 0; k < 16; k++) st_async_u32(ra_part + (par * kN + k * 64) * 4, y[k], rb_part + 8 * par);
            } else {
#pragma unroll
                for (int k = 0; k < 16; k++) dst_part[k * 64] = y[k];
            }
        }
        if (!ASYNC) cluster.sync();
        // ---- B: team j sums the two accumulator polynomials' contributions, inverse transform ----
        const uint32_t* rcv = ASYNC ? recv + par * kN : recv;
        const uint32_t* zrc = ASYNC ? zrecv + par * kN : zrecv;
        if (d == j) {
            if (ASYNC) mbar_wait_cluster(smem_u32(mbar + par), phase);
#pragma unroll
            for (int k = 0; k < 16; k++) y[k] += rcv[k * 64 + T];                      // (0, 8p)
            inv16(y, bufA, bufB, s_tab, xa, T, q, p, bar, 64);
#pragma unroll
            for (int k = 0; k < 16; k++) {
                zloc[T + 64 * k] = y[k];
                if (ASYNC) st_async_u32(ra_z + (par * kN + 64 * k) * 4, y[k], rb_z + 8 * par);
                else dst_z[64 * k] = y[k];
            }
        }
        if (ASYNC) {
            __syncthreads();                                   // zloc
            mbar_wait_cluster(smem_u32(mbar + 2 + par), phase);
        } else {
            cluster.sync();
        }
        // ---- C: CRT of output j, local accumulator update (8 coefficients per thread) ------------
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int e = (int)threadIdx.x + kCl4Threads * k;
            const uint32_t s0 = q ? zrc[e] : zloc[e];
            const uint32_t s1 = q ? zloc[e] : zrc[e];
            acc[e] += crt_lift32(s0, s1, c_crt_p0inv.w, c_crt_p0inv.ws);
        }
        __syncthreads();                                       // rotated reads of the next step cross threads
    }

    if (ASYNC) cluster.sync();                                 // no CTA leaves while a peer may still address it
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
