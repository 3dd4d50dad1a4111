// k_blindrotate_cl4s.cuh -- lowest-latency blind rotation: one gate per CLUSTER of 4 CTAs x 128
// threads, on the signed team transforms of ntt16s.cuh (the arithmetic of k_blind_rotate_s).
// CTA (q, j) = (prime, accumulator polynomial); its two 64-thread teams are the two gadget digits d
// of acc_j, i.e. TRGSW rows r = 2 j + d.
//
// A CMUX step moves 8 KB per CTA through distributed shared memory (which sustains only ~20 B/clk
// per CTA, /opt/skills/guides/B300_MICROARCH.md "CGA; DSMEM"), as 128-bit st.async stores that
// complete on an mbarrier of the receiving CTA:
//   A. team d: digit d of acc_j, forward transform (ten butterfly stages on the digits themselves: a
//      single polynomial per team does not amortise the table stages of k_blind_rotate_s).  The teams
//      swap their transformed digits through local shared memory and team d forms the partial sum of
//      OUTPUT d over both rows of acc_j in one Montgomery reduction (smontmul2).  The partial of
//      output j stays in registers (team j); the partial of output 1-j goes to CTA (q, 1-j)  [4 KB]
//   B. team j: adds the partial received from CTA (q, 1-j), inverse transform, result to CTA (1-q, j)
//      and to local memory                                                                  [4 KB]
//   C. all 128 threads: centred CRT of the two primes' results, acc_j += (CTAs (0, j) and (1, j)
//      keep identical copies of acc_j).
// Receive buffers are double-buffered by step parity: a sender is never two steps ahead of its
// receiver (its own progress needs data the receiver only sends after consuming the previous buffer).
// Lazy-residue bounds: forward < 12.5 p, partial sums < 2.56 p, inverse < 14.6 p (limit 16 p), the
// interval propagation of tests/emul/emul_ntt16s.cpp with these inputs.
//
// Replaces (for batches of a few gates) __BlindRotatePreAdd__ + Accumulate
// (reference include/gatebootstrapping_gpu.cuh:115-345), SampleExtract src/bootstrap_gpu.cu:366-381.
#pragma once
#include "k_ntt_s.cuh"
#ifdef B2_HOST_TRACE
namespace b2 {
uint32_t trace_mapa(uint32_t addr, int rank);
void trace_st_async(uint32_t remote_addr, int32_t a, int32_t b, int32_t c, int32_t d, uint32_t remote_bar);
int trace_cluster_rank();
void trace_cluster_sync();
}
namespace cooperative_groups {
struct cluster_group { unsigned block_rank() const { return (unsigned)b2::trace_cluster_rank(); } void sync() const { b2::trace_cluster_sync(); } };
inline cluster_group this_cluster() { return cluster_group{}; }
}
#else
#include <cooperative_groups.h>
#endif

namespace b2 {

#ifdef B2_HOST_TRACE
inline uint32_t mapa_u32(uint32_t addr, int rank) { return trace_mapa(addr, rank); }
inline void st_async_v4(uint32_t remote_addr, int32_t a, int32_t b, int32_t c, int32_t d, uint32_t remote_bar) { trace_st_async(remote_addr, a, b, c, d, remote_bar); }
#else
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, int rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
// 128-bit store into another CTA's shared memory that signals `bytes written` on that CTA's mbarrier
__device__ __forceinline__ void st_async_v4(uint32_t remote_addr, int32_t a, int32_t b, int32_t c, int32_t d, uint32_t remote_bar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.s32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(remote_addr),
                 "r"(a), "r"(b), "r"(c), "r"(d), "r"(remote_bar) : "memory");
}
#endif

#ifndef B2_CL4_STEPS
#define B2_CL4_STEPS kLwe
#endif
constexpr int kCl4Steps = B2_CL4_STEPS;     // 636; the host trace build (tests/emul/emul_cl4s.cpp) walks a few steps only
constexpr int kCl4Ctas = 4;
constexpr int kCl4Threads = 128;
constexpr int kCl4TabWords = (int)sizeof(PrimeSTables) / 4;
// my prime's tables | 2 teams x 2 exchange arrays | acc_j | the teams' transformed digits | received partial
// [parity] | my inverse result | the other prime's inverse result [parity] | abar | 4 mbarriers
constexpr int kCl4SmemWords = kCl4TabWords + 4 * kXsWords + kN + 2 * kN + 2 * kN + kN + 2 * kN + 640 / 2 + 8;
// 74 KB: three CTAs (of different gates, or of one gate when few are in flight) fit one SM
constexpr int kCl4SmemBytes = kCl4SmemWords * 4;

// JIT: schedule fuzzing (b2fhe_set_br_variant(6)): random sleeps around every hand-over, barrier and
// mbarrier wait; results must not change
// MUT: mutation test of that fuzzing (tools/fuzz_mutants.cu): the in-loop synchronisation point number MUT (1..5) is dropped
template <bool JIT = false, int MUT = 0>
__global__ void __cluster_dims__(kCl4Ctas, 1, 1) __launch_bounds__(kCl4Threads, 2)
    k_blind_rotate_cl4s(const BrParams P, const PrimeSTables* __restrict__ g_tab)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    B2_DYN_SMEM(smem);
    uint32_t* acc = smem + kCl4TabWords + 4 * kXsWords;        // acc_j
    uint32_t* tx = acc + kN;                                   // [team][v * 64 + T] as 128-bit vectors: transformed digit
    uint32_t* recv = tx + 2 * kN;                              // [parity][v * 64 + T]: partial sum of output j from CTA (q, 1-j)
    uint32_t* zloc = recv + 2 * kN;                            // [v * 64 + T]: inverse result, my prime (coefficients T + 64 (4 v + e))
    uint32_t* zrecv = zloc + kN;                               // [parity][v * 64 + T]: inverse result, other prime
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(zrecv + 2 * kN);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(zrecv + 2 * kN + 640 / 2);      // [0..1] recv, [2..3] zrecv

    const int d = threadIdx.x >> 6, T = threadIdx.x & 63;      // team = gadget digit
    const int rank = (int)cluster.block_rank();
    const int q = rank >> 1, j = rank & 1;
    const int item = blockIdx.x / kCl4Ctas;
    stage_prime_tables<1>(smem, g_tab + q);
    const SAddr a = team_addr(smem, 1, d, 0, T);
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
    for (int c = threadIdx.x; c < kN; c += kCl4Threads) B2_SMEM_WR(acc + c, j ? testvec_b((uint32_t)c, bbar, P.mu) : 0u);
    __syncthreads();

    const int32_t p = prime_of(q);
    const uint32_t pinv = pinv_of(q);
    // key vectors of OUTPUT d in the two rows of acc_j: [i][row][prime][out][256 x uint4]
    const uint4* bk_own = reinterpret_cast<const uint4*>(P.bk) + T + (size_t)(((2 * j + d) * 2 + q) * 2 + d) * (kN / 4);
    const uint4* bk_oth = reinterpret_cast<const uint4*>(P.bk) + T + (size_t)(((2 * j + 1 - d) * 2 + q) * 2 + d) * (kN / 4);
    constexpr size_t kStepVecs = (size_t)kRows * 4 * (kN / 4);
    const int dshift = d ? 16 : 24;
    // remote destinations: the partial of output 1-j -> CTA (q, 1-j); my inverse result -> CTA (1-q, j)
    const uint32_t ra_part = mapa_u32(smem_u32(recv + 4 * T), q * 2 + (1 - j));
    const uint32_t ra_z = mapa_u32(smem_u32(zrecv + 4 * T), (1 - q) * 2 + j);
    const uint32_t rb_part = mapa_u32(smem_u32(mbar), q * 2 + (1 - j));           // + 8 * parity
    const uint32_t rb_z = mapa_u32(smem_u32(mbar + 2), (1 - q) * 2 + j);
    const uint32_t tx_own = smem_u32(tx + d * kN + 4 * T), tx_oth = smem_u32(tx + (1 - d) * kN + 4 * T);
    const uint32_t zl = smem_u32(zloc + 4 * T);
    const int32_t cw = c_s.crt.w, cws = c_s.crt.ws;
    if (threadIdx.x == 0) {
        for (int b = 0; b < 4; b++) mbar_init(smem_u32(mbar + b), 1);
        mbar_fence_init();
    }
    cluster.sync();                                            // every CTA's shared memory (and mbarriers) are live

    uint4 kb[8];                                               // this step's key vectors, fetched ONE STEP AHEAD (L2 latency ~ 1/5 of a step)
#pragma unroll
    for (int v = 0; v < 4; v++) { kb[v] = __ldg(bk_own + v * 64); kb[4 + v] = __ldg(bk_oth + v * 64); }
    for (int i = 0; i < kCl4Steps; i++) {
        const uint32_t abar = s_abar[i];
        const uint32_t am = abar & (kN - 1);
        const bool flip = (abar >> kNbit) != 0;
        const uint32_t rbase = ((uint32_t)T - abar) & (kN - 1);
        const int kth = ((int)am - T + 63) >> 6;
        const int par = i & 1;
        const uint32_t phase = (uint32_t)(i >> 1) & 1u;
        if (threadIdx.x == 0) {                                // arm this step's two receive barriers: 4 KB each
            mbar_arrive_expect_tx(smem_u32(mbar + par), kN * 4);
            mbar_arrive_expect_tx(smem_u32(mbar + 2 + par), kN * 4);
        }
        // ---- A: digit polynomial, forward transform, pair products ------------------------------
        if (JIT) jitter_sleep(rank * 4 + (threadIdx.x >> 5));
        int32_t x[16], y[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            uint32_t rot = B2_SMEM_RD(acc + ((rbase + 64 * k) & (kN - 1)));
            const bool neg = (k < kth) != flip;
            rot = neg ? 0u - rot : rot;
            const uint32_t t = rot - B2_SMEM_RD(acc + T + 64 * k) + kDecompAdd;
            x[k] = (int32_t)((t >> dshift) & 0xFFu) - 128;
        }
        sfwd_general_dev<1>(x, y, a, q, p, bar, true, T);
        sts128<0>(tx_own, x[0], x[1], x[2], x[3]);
        sts128<1024>(tx_own, x[4], x[5], x[6], x[7]);
        sts128<2048>(tx_own, x[8], x[9], x[10], x[11]);
        sts128<3072>(tx_own, x[12], x[13], x[14], x[15]);
        if (JIT) jitter_sleep(0x11u + rank * 4 + (threadIdx.x >> 5));
        if (MUT != 1) __syncthreads();
        if (JIT) jitter_sleep(0x22u + rank * 4 + (threadIdx.x >> 5));
        {
            const V4 o0 = lds128<0>(tx_oth), o1 = lds128<1024>(tx_oth), o2 = lds128<2048>(tx_oth), o3 = lds128<3072>(tx_oth);
            const int32_t ox[16] = {o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w, o2.x, o2.y, o2.z, o2.w, o3.x, o3.y, o3.z, o3.w};
#pragma unroll
            for (int v = 0; v < 4; v++) {
                const uint4 km = kb[v], ko = kb[4 + v];
                y[4 * v + 0] = smontmul2(x[4 * v + 0], km.x, ox[4 * v + 0], ko.x, p, pinv);
                y[4 * v + 1] = smontmul2(x[4 * v + 1], km.y, ox[4 * v + 1], ko.y, p, pinv);
                y[4 * v + 2] = smontmul2(x[4 * v + 2], km.z, ox[4 * v + 2], ko.z, p, pinv);
                y[4 * v + 3] = smontmul2(x[4 * v + 3], km.w, ox[4 * v + 3], ko.w, p, pinv);
            }
        }
        {   // the products above were the last use of this step's key vectors: fetch the next ones now
            const size_t nxt = (size_t)(i + 1 < kLwe ? i + 1 : i) * kStepVecs;
#pragma unroll
            for (int v = 0; v < 4; v++) { kb[v] = __ldg(bk_own + nxt + v * 64); kb[4 + v] = __ldg(bk_oth + nxt + v * 64); }
        }
        if (JIT) jitter_sleep(0x33u + rank * 4 + (threadIdx.x >> 5));
        if (d != j) {                                          // output 1-j is finished by CTA (q, 1-j)
#pragma unroll
            for (int v = 0; v < 4; v++)
                st_async_v4(ra_part + (par * kN + v * 256) * 4, y[4 * v], y[4 * v + 1], y[4 * v + 2], y[4 * v + 3], rb_part + 8 * par);
        } else {
            // ---- B: team j adds the other accumulator polynomial's contribution, inverse transform ----
            if (MUT != 2) mbar_wait(smem_u32(mbar + par), phase);
            if (JIT) jitter_sleep(0x44u + rank * 4 + (threadIdx.x >> 5));
            const uint32_t rc = smem_u32(recv + par * kN + 4 * T);
            const V4 r0 = lds128<0>(rc), r1 = lds128<1024>(rc), r2 = lds128<2048>(rc), r3 = lds128<3072>(rc);
            const int32_t rv[16] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w, r3.x, r3.y, r3.z, r3.w};
#pragma unroll
            for (int k = 0; k < 16; k++) y[k] += rv[k];
            sinv1_dev(y, a, q, p, bar, T);
#pragma unroll
            for (int v = 0; v < 4; v++) {
                st_async_v4(ra_z + (par * kN + v * 256) * 4, y[4 * v], y[4 * v + 1], y[4 * v + 2], y[4 * v + 3], rb_z + 8 * par);
                if (v == 0) sts128<0>(zl, y[0], y[1], y[2], y[3]);
                if (v == 1) sts128<1024>(zl, y[4], y[5], y[6], y[7]);
                if (v == 2) sts128<2048>(zl, y[8], y[9], y[10], y[11]);
                if (v == 3) sts128<3072>(zl, y[12], y[13], y[14], y[15]);
            }
        }
        if (JIT) jitter_sleep(0x55u + rank * 4 + (threadIdx.x >> 5));
        if (MUT != 3) __syncthreads();                         // zloc
        if (MUT != 4) mbar_wait(smem_u32(mbar + 2 + par), phase);
        if (JIT) jitter_sleep(0x66u + rank * 4 + (threadIdx.x >> 5));
        // ---- C: CRT of output j, local accumulator update (two 128-bit slots per thread) ----------
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int slot = (int)threadIdx.x + kCl4Threads * h;                 // = v * 64 + T'
            const V4 mine = lds128<0>(smem_u32(zloc + 4 * slot));
            const V4 othr = lds128<0>(smem_u32(zrecv + par * kN + 4 * slot));
            const int32_t m[4] = {mine.x, mine.y, mine.z, mine.w}, o[4] = {othr.x, othr.y, othr.z, othr.w};
            uint32_t* ap = acc + (slot & 63) + 256 * (slot >> 6);                // coefficient T' + 64 (4 v + e)
#pragma unroll
            for (int e = 0; e < 4; e++) B2_SMEM_WR(ap + 64 * e, B2_SMEM_RD(ap + 64 * e) + (q ? scrt_lift32(o[e], m[e], cw, cws) : scrt_lift32(m[e], o[e], cw, cws)));
        }
        if (JIT) jitter_sleep(0x77u + rank * 4 + (threadIdx.x >> 5));
        if (MUT != 5) __syncthreads();                         // rotated reads of the next step cross threads
    }

    cluster.sync();                                            // no CTA leaves while a peer may still address it
    // ---- tail: CTA (0, 0) holds acc_a, CTA (0, 1) holds acc_b ----
    if (q != 0) return;
    if (P.trlwe) {
        uint32_t* out = P.out + (size_t)it.slot * 2 * kN + j * kN;
        for (int m = threadIdx.x; m < kN; m += kCl4Threads) out[m] = B2_SMEM_RD(acc + m);
        return;
    }
    uint32_t* out = P.out + (size_t)it.slot * kLvl1Words;      // SampleExtract(0), src/bootstrap_gpu.cu:366-381
    if (j == 0) { for (int m = threadIdx.x; m < kN; m += kCl4Threads) out[m] = m == 0 ? B2_SMEM_RD(acc) : 0u - B2_SMEM_RD(acc + kN - m); }
    else if (threadIdx.x == 0) out[kN] = B2_SMEM_RD(acc);
}

}  // namespace b2
