// k_polymul16.cuh -- batched negacyclic product mod 2^32 (BASELINE config 2,
// test/test_polynomial_mult_1024.cu:76-99 shape) on the 64-thread x 16-coefficient teams of
// ntt16.cuh: a product is 128 threads (one team per prime), G products per CTA.  Per team:
// two forward transforms (a reduced mod p, b small signed), a pointwise Montgomery product,
// one inverse transform; the teams then swap halves of their residues and each finishes the
// centred CRT of 512 coefficients.  Same arithmetic as k_polymul (k_ntt.cuh): identical bits.
#pragma once
#include "k_blindrotate16.cuh"

namespace b2 {

constexpr int kPolymulItemSmemWords = 4 * kXbufWords;
__host__ __device__ constexpr int polymul16_smem_bytes(int items) { return (kLane16TableWords + items * kPolymulItemSmemWords) * 4; }

template <int G>
__global__ void __launch_bounds__(128 * G, 1)
    k_polymul16(const uint32_t* __restrict__ a, const uint32_t* __restrict__ b, uint32_t* __restrict__ out, int npoly,
                const Lane16Tables* __restrict__ tab16)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    stage_tables16(s_tab, tab16);
    __syncthreads();
    const int gs = threadIdx.x >> 7, t128 = threadIdx.x & 127;
    const int q = t128 >> 6, T = t128 & 63;
    uint32_t* gbase = smem + kLane16TableWords + gs * kPolymulItemSmemWords;
    uint32_t* bufA = gbase + q * 2 * kXbufWords;
    uint32_t* bufB = bufA + kXbufWords;
    const uint32_t* bufA_other = gbase + (q ^ 1) * 2 * kXbufWords;
    const int bar = 1 + gs;
    const uint32_t p = q ? Prime<1>::p : Prime<0>::p;
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    const uint32_t r2 = c_r2[q];
    const X16 xa = make_x16(T);

    for (int pl = blockIdx.x * G + gs; pl < npoly; pl += gridDim.x * G) {      // whole teams loop together
        const uint32_t* pa = a + (size_t)pl * kN + T;
        const uint32_t* pb = b + (size_t)pl * kN + T;
        uint32_t x[16], y[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            x[k] = __ldg(pa + 64 * k) % p;                  // arbitrary u32 -> [0, p)
            y[k] = __ldg(pb + 64 * k) + p;                  // small signed -> [0, 2p)
        }
        fwd16_2(x, y, bufA, bufB, s_tab, xa, T, q, p, bar, false);        // both operands together: 16 butterflies in flight
        named_bar_sync(bar, 128);                           // its last loads were from both buffers; inv16 stores to bufA
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const int32_t m = umontmul_p(x[k], r2, p, pinv);                    // x R in (-p, p)
            const uint32_t xm = (uint32_t)(m + ((m >> 31) & (int32_t)p));       // canonical
            x[k] = (uint32_t)(umontmul_p(y[k], xm, p, pinv) + (int32_t)p);      // x y in (0, 2p)
        }
        inv16(x, bufA, bufB, s_tab, xa, T, q, p, bar);
        // layout A again (e = T + 64 k): team 0 finishes k < 8, team 1 the rest
        const int give = q == 0 ? 8 : 0, keep = 8 - give;
#pragma unroll
        for (int k = 0; k < 8; k++) bufA[T + 64 * k] = q == 0 ? x[8 + k] : x[k];
        named_bar_sync(bar, 128);
        uint32_t* po = out + (size_t)pl * kN + T;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t other = bufA_other[T + 64 * k];
            const uint32_t mine = q == 0 ? x[k] : x[8 + k];
            const uint32_t s0 = q == 0 ? mine : other;
            const uint32_t s1 = q == 0 ? other : mine;
            po[64 * (keep + k)] = crt_lift32(s0, s1, c_crt_p0inv.w, c_crt_p0inv.ws);
        }
        (void)give;
        named_bar_sync(bar, 128);                           // bufA is reused by the next product
    }
}

}  // namespace b2
