// k_ntt.cuh -- batched transform kernels: bootstrapping-key conversion and the
// NTT / INTT / negacyclic product microbenchmark entry points (BASELINE config 2).
//
// Replaces: __TRGSW2NTT__ (reference src/bootstrap_gpu.cu:43-70) and the test-only
// ForwardNTT / PointwiseMultiply / InverseNTT kernels
// (test/test_polynomial_mult_1024.cu:76-99).
//
// Mapping: one warp = one (polynomial, prime); a CTA of 256 threads = 4 polynomials.
// Shared memory: lane tables (24064 B) + 8 x 4 KiB exchange buffers.
#pragma once
#include "dev_common.cuh"

namespace b2 {

constexpr int kNttPolysPerCta = 4;
constexpr int kNttThreads = 64 * kNttPolysPerCta;
constexpr int kNttSmemBytes = kLaneTableWords * 4 + 2 * kNttPolysPerCta * 4096;

// residue in [0, p) of an unsigned 32-bit value (setup paths only; the compiler turns
// the constant division into a multiply-high sequence)
template <int Q>
__device__ __forceinline__ uint32_t reduce_u32(uint32_t v)
{
    return v % Prime<Q>::p;
}

// mode 0: generic polynomials -> canonical residues at bk_slot() order, out[poly][2][1024]
// mode 1: bootstrapping key   -> Montgomery-form residues, out index
//         ((pl>>1)*2 + Q)*2 + (pl&1), i.e. [i][row][prime][out poly][1024]
template <int Q>
__device__ __forceinline__ void fwd_store(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int pl,
                                          int mode, int signed_in, uint32_t* xbuf, const uint32_t* s_tab, int lane)
{
    uint32_t x[32];
    const uint32_t* src = in + (size_t)pl * kN;
#pragma unroll
    for (int k = 0; k < 32; k++) {
        const uint32_t v = __ldg(src + lane + 32 * k);
        x[k] = signed_in ? v + Prime<Q>::p : reduce_u32<Q>(v);      // small signed -> [0, 2p)
    }
    fwd_dev<Q>(x, xbuf, s_tab, lane);
    uint32_t* dst = mode == 1 ? out + ((size_t)((pl >> 1) * 2 + Q) * 2 + (pl & 1)) * kN
                              : out + ((size_t)pl * 2 + Q) * kN;
#pragma unroll
    for (int v = 0; v < 8; v++) {
        uint32_t r[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            // mode 1: x * R (Montgomery form); mode 0: x * 1 (plain), both canonical
            const uint32_t m = canon<Q>(umontmul<Q>(x[4 * v + e], c_r2[Q]));
            r[e] = mode == 1 ? m : canon<Q>(umontmul<Q>(m, 1u));
        }
        reinterpret_cast<uint4*>(dst)[v * 32 + lane] = make_uint4(r[0], r[1], r[2], r[3]);
    }
}

__global__ void __launch_bounds__(kNttThreads) k_ntt_fwd(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                         int npoly, int mode, int signed_in, const LaneTables* g_tab)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    stage_lane_tables(s_tab, g_tab);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* xbuf = smem + kLaneTableWords + warp * 1024;
    const int q = warp & 1;
    for (int pl = blockIdx.x * kNttPolysPerCta + (warp >> 1); pl < npoly; pl += gridDim.x * kNttPolysPerCta) {
        if (q == 0) fwd_store<0>(in, out, pl, mode, signed_in, xbuf, s_tab, lane);
        else fwd_store<1>(in, out, pl, mode, signed_in, xbuf, s_tab, lane);
    }
}

// inverse of mode-0 output: in[poly][2][1024] canonical residues -> out[poly][1024]
// = centred CRT value mod 2^32.  Also the tail of the product kernel.
template <int Q>
__device__ __forceinline__ void inv_crt_store(uint32_t (&x)[32], uint32_t* __restrict__ out, uint32_t* xbuf,
                                              uint32_t* xbuf_other, const uint32_t* s_tab, int lane, int barid)
{
    inv_dev<Q>(x, xbuf, s_tab, lane);
    __syncwarp();
    // warp 0 (prime 0) finishes coefficients k < 16, warp 1 the rest: swap halves
    constexpr int give = Q == 0 ? 16 : 0;      // the half I hand to my partner
#pragma unroll
    for (int k = 0; k < 16; k++) xbuf[(give + k) * 32 + lane] = x[give + k];
    named_bar_sync(barid, 64);
    constexpr int keep = 16 - give;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const uint32_t other = xbuf_other[(keep + k) * 32 + lane];
        const uint32_t s0 = Q == 0 ? x[keep + k] : other;
        const uint32_t s1 = Q == 0 ? other : x[keep + k];
        out[lane + 32 * (keep + k)] = crt_lift32(s0, s1, c_crt_p0inv.w, c_crt_p0inv.ws);
    }
    named_bar_sync(barid, 64);
}

__global__ void __launch_bounds__(kNttThreads) k_ntt_inv(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                         int npoly, const LaneTables* g_tab)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    stage_lane_tables(s_tab, g_tab);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* xbuf = smem + kLaneTableWords + warp * 1024;
    uint32_t* xbuf_other = smem + kLaneTableWords + (warp ^ 1) * 1024;
    const int q = warp & 1, barid = 1 + (warp >> 1);
    for (int pl = blockIdx.x * kNttPolysPerCta + (warp >> 1); pl < npoly; pl += gridDim.x * kNttPolysPerCta) {
        uint32_t x[32];
        const uint4* src = reinterpret_cast<const uint4*>(in + ((size_t)pl * 2 + q) * kN);
#pragma unroll
        for (int v = 0; v < 8; v++) {
            const uint4 t = __ldg(src + v * 32 + lane);
            x[4 * v] = t.x; x[4 * v + 1] = t.y; x[4 * v + 2] = t.z; x[4 * v + 3] = t.w;
        }
        if (q == 0) inv_crt_store<0>(x, out + (size_t)pl * kN, xbuf, xbuf_other, s_tab, lane, barid);
        else inv_crt_store<1>(x, out + (size_t)pl * kN, xbuf, xbuf_other, s_tab, lane, barid);
    }
}

// out = a (*) b negacyclic, mod 2^32.  a: arbitrary u32, b: signed small (|b| <= 2^10
// keeps |sum| < 2^52; see b2fhe.h for the exactness domain).
template <int Q>
__device__ __forceinline__ void polymul_body(const uint32_t* __restrict__ a, const uint32_t* __restrict__ b,
                                             uint32_t* __restrict__ out, uint32_t* xbuf, uint32_t* xbuf_other,
                                             const uint32_t* s_tab, int lane, int barid)
{
    uint32_t x[32], y[32];
#pragma unroll
    for (int k = 0; k < 32; k++) {
        x[k] = reduce_u32<Q>(__ldg(a + lane + 32 * k));
        y[k] = __ldg(b + lane + 32 * k) + Prime<Q>::p;            // small signed -> [0, 2p)
    }
    fwd_dev<Q>(x, xbuf, s_tab, lane);
    fwd_dev<Q>(y, xbuf, s_tab, lane);
#pragma unroll
    for (int k = 0; k < 32; k++) {
        const uint32_t xm = canon<Q>(umontmul<Q>(x[k], c_r2[Q]));              // x * R in [0, p)
        x[k] = (uint32_t)(umontmul<Q>(y[k], xm) + (int32_t)Prime<Q>::p);       // x * y in (0, 2p)
    }
    inv_crt_store<Q>(x, out, xbuf, xbuf_other, s_tab, lane, barid);
}

__global__ void __launch_bounds__(kNttThreads) k_polymul(const uint32_t* __restrict__ a, const uint32_t* __restrict__ b,
                                                         uint32_t* __restrict__ out, int npoly, const LaneTables* g_tab)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    stage_lane_tables(s_tab, g_tab);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* xbuf = smem + kLaneTableWords + warp * 1024;
    uint32_t* xbuf_other = smem + kLaneTableWords + (warp ^ 1) * 1024;
    const int q = warp & 1, barid = 1 + (warp >> 1);
    for (int pl = blockIdx.x * kNttPolysPerCta + (warp >> 1); pl < npoly; pl += gridDim.x * kNttPolysPerCta) {
        const size_t o = (size_t)pl * kN;
        if (q == 0) polymul_body<0>(a + o, b + o, out + o, xbuf, xbuf_other, s_tab, lane, barid);
        else polymul_body<1>(a + o, b + o, out + o, xbuf, xbuf_other, s_tab, lane, barid);
    }
}

// ---- integer-pipe peak probe (roofline denominator) -------------------
// mode 0: independent IMAD (lo) chains; mode 1: the butterfly actually used
// (IMAD.HI + 2 IMAD + IADD3 + IADD), 16 independent chains per thread, registers only.
__global__ void __launch_bounds__(256) k_int_peak(uint32_t* out, int iters, int mode, uint32_t seed)
{
    uint32_t x[16];
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = seed * (threadIdx.x + 1) + k * 0x9E3779B9u;
    const uint32_t w = (seed | 1u) % Prime<0>::p;
    const Tw tw{w, seed * 2654435761u};
    if (mode == 0) {
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int k = 0; k < 16; k++) x[k] = x[k] * w + seed;
        }
    } else {
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int k = 0; k < 16; k += 2) butterfly<0>(x[k], x[k + 1], tw);
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) acc ^= x[k];
    if (acc == 0x12345678u) out[0] = acc;      // keeps the loop alive, practically never taken
}

}  // namespace b2
