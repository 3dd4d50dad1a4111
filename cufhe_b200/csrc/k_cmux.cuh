// k_cmux.cuh -- batched CMUX with a caller-supplied NTT-domain TRGSW:
//     out = c0 + (c1 - c0) [.] TRGSW            (TRLWE<lvl1param> operands)
// Replaces __CMUXNTT__ / TRLWESubAndDecomposition and __TRGSW2NTT__ (reference
// src/bootstrap_gpu.cu:43-70,162-285; API src/cufhe_gates_gpu.cu:69-85), which run one
// CMUX per 512-thread block.  Here an item is 128 threads (two 64-thread teams, one per
// prime, 16 coefficients per thread: the external-product step of k_blind_rotate_s with the
// rotation replaced by c1 - c0 and the digits transformed by the general butterfly passes), G
// items per CTA.
//
// NTT-domain TRGSW format (this engine's, produced by b2fhe_trgsw2ntt_batch): per TRGSW
// [4 rows][2 primes][2 outputs][1024] u32 Montgomery residues in bk16_slot order (65536 B)
// -- the layout of one bootstrapping-key row.
#pragma once
#include "k_ntt_s.cuh"

namespace b2 {

struct CmuxParams {
    uint32_t* out;            // [count][2][1024]
    const uint32_t* cs;       // NTT-domain TRGSWs
    const uint32_t* c1;       // [count][2][1024]
    const uint32_t* c0;       // [count][2][1024]
    int count;
    int cs_per_item;          // 1: item i uses TRGSW i; 0: all items share TRGSW 0
};

// two primes' tables | per item 2 teams x 2 exchange arrays
__host__ __device__ constexpr int cmux_smem_bytes(int items) { return 2 * (int)sizeof(PrimeSTables) + items * 2 * 2 * kXsBytes; }
static_assert(cmux_smem_bytes(4) <= 227 * 1024, "shared memory");

template <int G>
__global__ void __launch_bounds__(128 * G, 1) k_cmux_s(const CmuxParams P, const PrimeSTables* __restrict__ g_tab)
{
    B2_DYN_SMEM(smem);
    stage_prime_tables<2>(smem, g_tab);
    __syncthreads();
    const int gs = threadIdx.x >> 7, t128 = threadIdx.x & 127;
    const int q = t128 >> 6, T = t128 & 63;
    const int item = blockIdx.x * G + gs;
    if (item >= P.count) return;
    const SAddr a = team_addr(smem, 2, 2 * gs + q, q, T);
    const uint32_t own_other = team_addr(smem, 2, 2 * gs + (q ^ 1), q ^ 1, T).own;
    const int tbar = 1 + 2 * gs + q, pbar = 1 + 2 * G + gs;       // my team (64 threads), my item (128)

    const int32_t p = prime_of(q);
    const uint32_t pinv = pinv_of(q);
    const uint32_t* c1 = P.c1 + (size_t)item * 2 * kN;
    const uint32_t* c0 = P.c0 + (size_t)item * 2 * kN;
    const uint4* cs4 = reinterpret_cast<const uint4*>(P.cs) + (size_t)(P.cs_per_item ? item : 0) * (kRows * 4 * (kN / 4)) + T +
                       (size_t)q * 2 * (kN / 4);
    const int off_a = q * (kN / 4), off_b = (1 - q) * (kN / 4);

    // the two digits of c1_j - c0_j are transformed together and share one Montgomery reduction per output;
    // sa = output q (this team finishes it), sb = output 1-q: |.| < 2 (2 * 12.5 / 32 + 0.5) p = 2.57 p
    int32_t sa[16], sb[16];
#pragma unroll
    for (int k = 0; k < 16; k++) sa[k] = sb[k] = 0;
#pragma unroll 1
    for (int j = 0; j < 2; j++) {
        int32_t x0[16], x1[16];
        const uint32_t* a1 = c1 + j * kN + T;
        const uint32_t* a0 = c0 + j * kN + T;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const uint32_t t = a1[64 * k] - a0[64 * k] + kDecompAdd;   // :162-194 (plain loads: out may alias)
            x0[k] = (int32_t)(t >> 24) - 128;
            x1[k] = (int32_t)((t >> 16) & 0xFFu) - 128;
        }
        sfwd_general_dev<2>(x0, x1, a, q, p, tbar, j != 0, T);
        const uint4* row0 = cs4 + (size_t)(2 * j) * (4 * (kN / 4));
        const uint4* row1 = row0 + 4 * (kN / 4);
#pragma unroll
        for (int v = 0; v < 4; v++) {
            const uint4 b0 = __ldg(row0 + off_a + v * 64), b1 = __ldg(row1 + off_a + v * 64);
            sa[4 * v + 0] += smontmul2(x0[4 * v + 0], b0.x, x1[4 * v + 0], b1.x, p, pinv);
            sa[4 * v + 1] += smontmul2(x0[4 * v + 1], b0.y, x1[4 * v + 1], b1.y, p, pinv);
            sa[4 * v + 2] += smontmul2(x0[4 * v + 2], b0.z, x1[4 * v + 2], b1.z, p, pinv);
            sa[4 * v + 3] += smontmul2(x0[4 * v + 3], b0.w, x1[4 * v + 3], b1.w, p, pinv);
        }
#pragma unroll
        for (int v = 0; v < 4; v++) {
            const uint4 b0 = __ldg(row0 + off_b + v * 64), b1 = __ldg(row1 + off_b + v * 64);
            sb[4 * v + 0] += smontmul2(x0[4 * v + 0], b0.x, x1[4 * v + 0], b1.x, p, pinv);
            sb[4 * v + 1] += smontmul2(x0[4 * v + 1], b0.y, x1[4 * v + 1], b1.y, p, pinv);
            sb[4 * v + 2] += smontmul2(x0[4 * v + 2], b0.z, x1[4 * v + 2], b1.z, p, pinv);
            sb[4 * v + 3] += smontmul2(x0[4 * v + 3], b0.w, x1[4 * v + 3], b1.w, p, pinv);
        }
    }
    sinv1_dev(sb, a, q, p, tbar, T);
    named_bar_sync(tbar, 64);            // the team has read the last exchange of that transform
    sinv1_dev(sa, a, q, p, tbar, T);
    named_bar_sync(tbar, 64);
    // sb = INTT(output 1-q) -> the other team, through chunk T of my first exchange array
    sts128<0>(a.own, sb[0], sb[1], sb[2], sb[3]); sts128<16>(a.own, sb[4], sb[5], sb[6], sb[7]);
    sts128<32>(a.own, sb[8], sb[9], sb[10], sb[11]); sts128<48>(a.own, sb[12], sb[13], sb[14], sb[15]);
    named_bar_sync(pbar, 128);
    int32_t ot[16];
    ld_own<0>(ot, own_other);
    uint32_t* out = P.out + (size_t)item * 2 * kN + q * kN + T;
    const uint32_t* base = c0 + q * kN + T;
    const int32_t cw = c_s.crt.w, cws = c_s.crt.ws;
#pragma unroll
    for (int k = 0; k < 16; k++)
        out[64 * k] = base[64 * k] + (q ? scrt_lift32(ot[k], sa[k], cw, cws) : scrt_lift32(sa[k], ot[k], cw, cws));
}

}  // namespace b2
