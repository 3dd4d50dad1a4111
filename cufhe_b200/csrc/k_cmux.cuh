// k_cmux.cuh -- batched CMUX with a caller-supplied NTT-domain TRGSW:
//     out = c0 + (c1 - c0) [.] TRGSW            (TRLWE<lvl1param> operands)
// Replaces __CMUXNTT__ / TRLWESubAndDecomposition and __TRGSW2NTT__ (reference
// src/bootstrap_gpu.cu:43-70,162-285; API src/cufhe_gates_gpu.cu:69-85), which run one
// CMUX per 512-thread block.  Here an item is 128 threads (two 64-thread teams, one per
// prime, 16 coefficients per thread: the external-product step of k_blind_rotate16 with the
// rotation replaced by c1 - c0), G items per CTA.
//
// NTT-domain TRGSW format (this engine's, produced by b2fhe_trgsw2ntt_batch): per TRGSW
// [4 rows][2 primes][2 outputs][1024] u32 Montgomery residues in bk16_slot order (65536 B)
// -- the layout of one bootstrapping-key row.
#pragma once
#include "team16_dev.cuh"

namespace b2 {

struct CmuxParams {
    uint32_t* out;            // [count][2][1024]
    const uint32_t* cs;       // NTT-domain TRGSWs
    const uint32_t* c1;       // [count][2][1024]
    const uint32_t* c0;       // [count][2][1024]
    int count;
    int cs_per_item;          // 1: item i uses TRGSW i; 0: all items share TRGSW 0
};

constexpr int kCmuxItemSmemWords = 4 * kXbufWords;
__host__ __device__ constexpr int cmux_smem_bytes(int items) { return (kLane16TableWords + items * kCmuxItemSmemWords) * 4; }

template <int G>
__global__ void __launch_bounds__(128 * G, 1) k_cmux16(const CmuxParams P, const Lane16Tables* __restrict__ tab16)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    stage_tables16(s_tab, tab16);
    __syncthreads();
    const int gs = threadIdx.x >> 7, t128 = threadIdx.x & 127;
    const int q = t128 >> 6, T = t128 & 63;
    const int item = blockIdx.x * G + gs;
    if (item >= P.count) return;
    uint32_t* gbase = smem + kLane16TableWords + gs * kCmuxItemSmemWords;
    uint32_t* bufA = gbase + q * 2 * kXbufWords;
    uint32_t* bufB = bufA + kXbufWords;
    const uint32_t* bufA_other = gbase + (q ^ 1) * 2 * kXbufWords;
    const int bar = 1 + gs;

    const uint32_t p = q ? Prime<1>::p : Prime<0>::p;
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    const X16 xa = make_x16(T);
    const uint32_t* c1 = P.c1 + (size_t)item * 2 * kN;
    const uint32_t* c0 = P.c0 + (size_t)item * 2 * kN;
    const uint4* cs4 = reinterpret_cast<const uint4*>(P.cs) + (size_t)(P.cs_per_item ? item : 0) * (kRows * 4 * (kN / 4)) + T +
                       (size_t)q * 2 * (kN / 4);
    const int off_a = q * (kN / 4), off_b = (1 - q) * (kN / 4);

    // same schedule as k_blind_rotate16's default: the two digits of c1_j - c0_j are transformed together,
    // a row pair shares one Montgomery reduction, the two inverse transforms run together
    uint32_t sa[16], sb[16];
#pragma unroll
    for (int k = 0; k < 16; k++) sa[k] = sb[k] = 2u * p;
#pragma unroll 1
    for (int j = 0; j < 2; j++) {
        uint32_t x0[16], x1[16];
        const uint32_t* a1 = c1 + j * kN + T;
        const uint32_t* a0 = c0 + j * kN + T;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const uint32_t t = a1[64 * k] - a0[64 * k] + kDecompAdd;   // :162-194 (plain loads: out may alias)
            x0[k] = t >> 24;                                           // raw digits: stand for x - 128
            x1[k] = (t >> 16) & 0xFFu;
        }
        fwd16_digits2(x0, x1, bufA, bufB, s_tab, xa, T, q, p, bar, j != 0);
        const uint4* row0 = cs4 + (size_t)(2 * j) * (4 * (kN / 4));
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
    named_bar_sync(bar, 128);            // the paired transform's last loads were from both buffers
    inv16_2(sb, sa, bufA, bufB, s_tab, xa, T, q, p, bar);
    named_bar_sync(bar, 128);            // ... and so were this one's; the hand-over goes through bufA
    // sa = INTT(output q) (this team finishes it), sb = INTT(output 1-q) -> other team
#pragma unroll
    for (int k = 0; k < 16; k++) bufA[T + 64 * k] = sb[k];
    named_bar_sync(bar, 128);
    uint32_t* out = P.out + (size_t)item * 2 * kN + q * kN + T;
    const uint32_t* base = c0 + q * kN + T;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const uint32_t t = bufA_other[T + 64 * k];
        const uint32_t s0 = q ? t : sa[k];
        const uint32_t s1 = q ? sa[k] : t;
        out[64 * k] = base[64 * k] + crt_lift32(s0, s1, c_crt_p0inv.w, c_crt_p0inv.ws);
    }
}

}  // namespace b2
