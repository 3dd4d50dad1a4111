// k_blindrotate16q.cuh -- k_blind_rotate16 with paired transforms and FOUR exchange buffers per team.
//
// Same mathematics, key layout, tables and bits as k_blind_rotate16 (replaces
// __BlindRotatePreAdd__/__BlindRotate__/Accumulate, reference
// include/gatebootstrapping_gpu.cuh:115-345, and __SampleExtractIndex__,
// src/bootstrap_gpu.cu:366-381).  What differs is the schedule of a CMUX step:
//   * the two gadget digits of an accumulator polynomial are transformed TOGETHER (16 independent
//     butterflies per stage per thread instead of 8), and so are the two inverse transforms;
//   * a row pair shares one Montgomery reduction (x0 b0 + x1 b1 summed in 64 bits by IMAD.WIDE
//     with accumulate): 7 multiply-pipe slots per pair instead of 10;
//   * a team owns four exchange buffers: the first exchange of every paired transform goes
//     through A and B, the second through C and D, so a buffer is never stored to before a
//     barrier has passed since its last load -- 8 gate barriers per step instead of 14.
// The fourth buffer fits (4 gates: 230,912 B of the 232,448 B a CTA may take) because the 636
// switched mask values are no longer staged in shared memory: the head parks them in the gate's
// own OUTPUT slot in global memory (not written until the tail) and the loop fetches one value a
// step ahead.
#pragma once
#include "k_blindrotate16.cuh"

namespace b2 {

constexpr int kBr16qGateSmemWords = 2 * kN + 8 * kXbufWords;     // acc + 2 teams x 4 buffers
__host__ __device__ constexpr int br16q_smem_bytes(int gates) { return (kLane16TableWords + gates * kBr16qGateSmemWords) * 4; }
static_assert(br16q_smem_bytes(4) <= 227 * 1024, "four gates per CTA must fit");

// buf = the team's four buffers A, B, C, D back to back
__device__ __forceinline__ void fwd16q_digits2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* buf, const uint32_t* s_tab,
                                               const X16& a, int T, int q, uint32_t p, int bar)
{
    uint32_t *bufA = buf, *bufB = buf + kXbufWords, *bufC = buf + 2 * kXbufWords, *bufD = buf + 3 * kXbufWords;
    pass16_fwd_digits(x0, t16_lut1(s_tab, q), t16_lut24(s_tab, q), t16_lut35(s_tab, q), c_fw1[q], p);
    pass16_fwd_digits(x1, t16_lut1(s_tab, q), t16_lut24(s_tab, q), t16_lut35(s_tab, q), c_fw1[q], p);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.wA + 68 * k] = x0[k]; bufB[a.wA + 68 * k] = x1[k]; }
    named_bar_sync(bar, 128);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[a.rB + 4 * k]; x1[k] = bufB[a.rB + 4 * k]; }
    pass16_fwd<4, 8>(x0, t16_fw2(s_tab, q, T >> 2), 16, p);
    pass16_fwd<4, 8>(x1, t16_fw2(s_tab, q, T >> 2), 16, p);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufC[a.wB + 4 * k + (k >> 2)] = x0[k]; bufD[a.wB + 4 * k + (k >> 2)] = x1[k]; }
    named_bar_sync(bar, 128);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufC[a.rC + k]; x1[k] = bufD[a.rC + k]; }
    pass16_fwd3(x0, t16_fw3(s_tab, q, T), 64, p);
    pass16_fwd3(x1, t16_fw3(s_tab, q, T), 64, p);
}

__device__ __forceinline__ void inv16q_2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* buf, const uint32_t* s_tab,
                                         const X16& a, int T, int q, uint32_t p, int bar)
{
    uint32_t *bufA = buf, *bufB = buf + kXbufWords, *bufC = buf + 2 * kXbufWords, *bufD = buf + 3 * kXbufWords;
    pass16_inv<true>(x0, c_iv1[q], 1, p);
    pass16_inv<true>(x1, c_iv1[q], 1, p);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.rC + k] = x0[k]; bufB[a.rC + k] = x1[k]; }
    named_bar_sync(bar, 128);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[a.rD + 17 * k]; x1[k] = bufB[a.rD + 17 * k]; }
    pass16_inv<false>(x0, t16_iv2(s_tab, q, T & 15), 16, p);
    pass16_inv<false>(x1, t16_iv2(s_tab, q, T & 15), 16, p);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufC[a.wD + 16 * k] = x0[k]; bufD[a.wD + 16 * k] = x1[k]; }
    named_bar_sync(bar, 128);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufC[T + 64 * k + 16 * (k >> 2)]; x1[k] = bufD[T + 64 * k + 16 * (k >> 2)]; }
    pass16_inv3(x0, t16_iv3(s_tab, q, T), 64, p);
    pass16_inv3(x1, t16_iv3(s_tab, q, T), 64, p);
    const Tw* sc = t16_scale(s_tab, q, T);
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const Tw f = sc[k * 64];
        x0[k] = shoup_mul_p(x0[k], f.w, f.ws, p);
        x1[k] = shoup_mul_p(x1[k], f.w, f.ws, p);
    }
}

// g_abar: the gate's 636 switched mask values in global memory (written by this gate's threads
// before its first barrier; read through L2).
__device__ __forceinline__ void br16q_gate_loop(const uint32_t* __restrict__ bk, uint32_t* acc, uint32_t* buf,
                                                const uint32_t* bufA_other, const uint16_t* g_abar,
                                                const uint32_t* s_tab, const int T, const int q, const int bar)
{
    const uint32_t p = q ? Prime<1>::p : Prime<0>::p;
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    const X16 xa = make_x16(T);
    const uint4* bk4 = reinterpret_cast<const uint4*>(bk) + T + (size_t)q * 2 * (kN / 4);
    const int off_a = q * (kN / 4), off_b = (1 - q) * (kN / 4);
    uint32_t* acc_mine = acc + q * kN + T;
    uint32_t abar_next = __ldcg(g_abar);
    for (int i = 0; i < kLwe; i++) {
        const uint32_t abar = abar_next;
        abar_next = __ldcg(g_abar + (i + 1 < kLwe ? i + 1 : i));
        uint32_t sa[16], sb[16];
#pragma unroll
        for (int k = 0; k < 16; k++) sa[k] = sb[k] = 2u * p;       // + two pair products in (-p, 2p) -> (0, 6p)

        const uint32_t am = abar & (kN - 1);
        const bool flip = (abar >> kNbit) != 0;
        const uint32_t rbase = ((uint32_t)T - abar) & (kN - 1);
        const int kth = ((int)am - T + 63) >> 6;       // coefficient T + 64k is negated iff (k < kth) xor flip
#pragma unroll 1
        for (int j = 0; j < 2; j++) {                  // accumulator polynomial j = TRGSW rows 2j, 2j+1
            uint32_t x0[16], x1[16];
            const uint32_t* aj = acc + j * kN;
#pragma unroll
            for (int k = 0; k < 16; k++) {
                uint32_t rot = aj[(rbase + 64 * k) & (kN - 1)];
                const bool neg = (k < kth) != flip;
                rot = neg ? 0u - rot : rot;
                const uint32_t t = rot - aj[T + 64 * k] + kDecompAdd;
                x0[k] = t >> 24;                       // raw digits: stand for x - 128
                x1[k] = (t >> 16) & 0xFFu;
            }
            fwd16q_digits2(x0, x1, buf, s_tab, xa, T, q, p, bar);
            const uint4* row0 = bk4 + (size_t)(i * kRows + 2 * j) * (4 * (kN / 4));
            const uint4* row1 = row0 + 4 * (kN / 4);
#pragma unroll
            for (int v = 0; v < 4; v++) {
                const uint4 b0 = __ldg(row0 + off_a + v * 64), b1 = __ldg(row1 + off_a + v * 64);
                sa[4 * v + 0] += umontmul2_p(x0[4 * v + 0], b0.x, x1[4 * v + 0], b1.x, p, pinv);
                sa[4 * v + 1] += umontmul2_p(x0[4 * v + 1], b0.y, x1[4 * v + 1], b1.y, p, pinv);
                sa[4 * v + 2] += umontmul2_p(x0[4 * v + 2], b0.z, x1[4 * v + 2], b1.z, p, pinv);
                sa[4 * v + 3] += umontmul2_p(x0[4 * v + 3], b0.w, x1[4 * v + 3], b1.w, p, pinv);
            }
#pragma unroll
            for (int v = 0; v < 4; v++) {
                const uint4 b0 = __ldg(row0 + off_b + v * 64), b1 = __ldg(row1 + off_b + v * 64);
                sb[4 * v + 0] += umontmul2_p(x0[4 * v + 0], b0.x, x1[4 * v + 0], b1.x, p, pinv);
                sb[4 * v + 1] += umontmul2_p(x0[4 * v + 1], b0.y, x1[4 * v + 1], b1.y, p, pinv);
                sb[4 * v + 2] += umontmul2_p(x0[4 * v + 2], b0.z, x1[4 * v + 2], b1.z, p, pinv);
                sb[4 * v + 3] += umontmul2_p(x0[4 * v + 3], b0.w, x1[4 * v + 3], b1.w, p, pinv);
            }
        }
        inv16q_2(sb, sa, buf, s_tab, xa, T, q, p, bar);
        // sa = INTT(out q) (mine), sb = INTT(out 1-q): hand sb to the other team through buffer A
        // (last loaded before the inverse transform's second barrier)
#pragma unroll
        for (int k = 0; k < 16; k++) buf[T + 64 * k] = sb[k];
        named_bar_sync(bar, 128);
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const uint32_t t = bufA_other[T + 64 * k];
            const uint32_t s0 = q ? t : sa[k];
            const uint32_t s1 = q ? sa[k] : t;
            acc_mine[64 * k] += crt_lift32(s0, s1, c_crt_p0inv.w, c_crt_p0inv.ws);
        }
        named_bar_sync(bar, 128);
    }
}

template <int G>
__global__ void __launch_bounds__(128 * G, 1) k_blind_rotate16q(const BrParams P, const Lane16Tables* __restrict__ tab16)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    stage_tables16(s_tab, tab16);
    __syncthreads();

    const int gs = threadIdx.x >> 7, t128 = threadIdx.x & 127;
    const int q = t128 >> 6, T = t128 & 63;
    const int item = blockIdx.x * G + gs;
    if (item >= P.count) return;

    uint32_t* gbase = smem + kLane16TableWords + gs * kBr16qGateSmemWords;
    uint32_t* acc = gbase;
    uint32_t* buf = gbase + 2 * kN + q * 4 * kXbufWords;
    const uint32_t* bufA_other = gbase + 2 * kN + (q ^ 1) * 4 * kXbufWords;
    const int bar = 1 + gs;

    // ---- pre-add + modulus switch (gatebootstrapping_gpu.cuh:316-345) ----
    BrItem it;
    if (P.items) it = P.items[item];
    else { it.g0 = it.g1 = it.slot = (uint32_t)item; it.ca = (int8_t)P.ca; it.cb = (int8_t)P.cb; it.src1 = 0; it.off = P.off; }
    uint32_t* out = P.out + (size_t)it.slot * (P.trlwe ? 2 * kN : kLvl1Words);
    uint16_t* g_abar = reinterpret_cast<uint16_t*>(out);     // parked in the output slot until the tail overwrites it
    const uint16_t* a0 = P.in0 + (size_t)it.g0 * kLvl0Words;
    const uint16_t* src1 = it.src1 ? P.in2 : P.in1;
    const uint16_t* a1 = (it.cb != 0 && src1) ? src1 + (size_t)it.g1 * kLvl0Words : nullptr;
    const int ca = it.ca, cb = a1 ? it.cb : 0;
    for (int i = t128; i < kLwe; i += 128) {
        const int c = ca * (int)a0[i] + (a1 ? cb * (int)a1[i] : 0);
        g_abar[i] = (uint16_t)modswitch_a((uint32_t)c & 0xFFFFu);
    }
    const int cbody = it.off + ca * (int)a0[kLwe] + (a1 ? cb * (int)a1[kLwe] : 0);
    const uint32_t bbar = modswitch_b((uint32_t)cbody & 0xFFFFu);
    for (int c = t128; c < kN; c += 128) {
        acc[c] = 0u;
        acc[kN + c] = testvec_b((uint32_t)c, bbar, P.mu);
    }
    __threadfence_block();
    named_bar_sync(bar, 128);

    br16q_gate_loop(P.bk, acc, buf, bufA_other, g_abar, s_tab, T, q, bar);

    // ---- SampleExtract(0) (src/bootstrap_gpu.cu:366-381) ----
    // (the closing barrier of the last step is behind every load of g_abar)
    if (P.trlwe) {          // BootstrapTLWE2TRLWE (src/bootstrap_gpu.cu:316-322): no extraction
        for (int m = t128; m < 2 * kN; m += 128) out[m] = acc[m];
        return;
    }
    for (int m = t128; m < kN; m += 128) out[m] = m == 0 ? acc[0] : 0u - acc[kN - m];
    if (t128 == 0) out[kN] = acc[kN];
}

}  // namespace b2
