// dev_common.cuh -- device-side plumbing shared by all kernels: constant twiddles,
// per-lane tables staged into shared memory, the warp transpose, full transforms.
#pragma once
#include <cuda_runtime.h>
#include "ntt_core.cuh"
#include "tfhe_core.cuh"

namespace b2 {

// warp-uniform twiddles: after unrolling every index is a compile-time constant, so
// ptxas folds them into c[bank][imm] operands of IMAD.WIDE (no load instruction).
__constant__ Tw c_fw1[2][32];
__constant__ Tw c_iv1[2][32];
__constant__ uint32_t c_r2[2];
__constant__ Tw c_crt_p0inv;

// per-lane tables (global copy; kernels stage them into shared memory once)
struct LaneTables {
    Tw fw2[2][31][32];
    Tw iv2[2][31][32];
    Tw scale[2][32][32];
};
constexpr int kLaneTableWords = sizeof(LaneTables) / 4;   // 12032 words = 48128 B

__device__ __forceinline__ void stage_lane_tables(uint32_t* s_tab, const LaneTables* g_tab)
{
    const uint4* src = reinterpret_cast<const uint4*>(g_tab);
    uint4* dst = reinterpret_cast<uint4*>(s_tab);
    for (int i = threadIdx.x; i < kLaneTableWords / 4; i += blockDim.x) dst[i] = __ldg(src + i);
}
__device__ __forceinline__ const Tw* tab_fw2(const uint32_t* s_tab, int q, int lane) { return reinterpret_cast<const Tw*>(s_tab) + q * 31 * 32 + lane; }
__device__ __forceinline__ const Tw* tab_iv2(const uint32_t* s_tab, int q, int lane) { return reinterpret_cast<const Tw*>(s_tab) + 2 * 31 * 32 + q * 31 * 32 + lane; }
__device__ __forceinline__ const Tw* tab_scale(const uint32_t* s_tab, int q, int lane) { return reinterpret_cast<const Tw*>(s_tab) + 4 * 31 * 32 + q * 32 * 32 + lane; }

// 32x32 transpose of a warp's register tile through its private 4 KiB buffer
__device__ __forceinline__ void exchange_dev(uint32_t (&x)[32], uint32_t* buf, int lane)
{
    __syncwarp();
#pragma unroll
    for (int v = 0; v < 8; v++)
        *reinterpret_cast<int4*>(buf + xchg_wr_word(lane, v)) = make_int4((int)x[4 * v], (int)x[4 * v + 1], (int)x[4 * v + 2], (int)x[4 * v + 3]);
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 32; j++) x[j] = buf[xchg_rd_word(lane, j)];
}

// forward transform: layout A in, layout B out
template <int Q>
__device__ __forceinline__ void fwd_dev(uint32_t (&x)[32], uint32_t* xbuf, const uint32_t* s_tab, int lane)
{
    pass_fwd<Q>(x, c_fw1[Q], 1);
    exchange_dev(x, xbuf, lane);
    pass_fwd<Q>(x, tab_fw2(s_tab, Q, lane), 32);
}
// inverse transform incl. n^-1 psi^-c scaling: layout B in, layout A out, values in [0, 2p)
template <int Q>
__device__ __forceinline__ void inv_dev(uint32_t (&x)[32], uint32_t* xbuf, const uint32_t* s_tab, int lane)
{
    pass_inv<Q>(x, c_iv1[Q], 1);
    exchange_dev(x, xbuf, lane);
    pass_inv<Q>(x, tab_iv2(s_tab, Q, lane), 32);
    const Tw* sc = tab_scale(s_tab, Q, lane);
#pragma unroll
    for (int k = 0; k < 32; k++) {
        const Tw f = sc[k * 32];
        x[k] = shoup_mul<Q>(x[k], f.w, f.ws);
    }
}

// run-time-prime forms (one code copy for both primes; q is warp-uniform)
__device__ __forceinline__ void fwd_dev_rt(uint32_t (&x)[32], uint32_t* xbuf, const uint32_t* s_tab, int lane, int q, uint32_t p)
{
    pass_fwd_p(x, c_fw1[q], 1, p);
    exchange_dev(x, xbuf, lane);
    pass_fwd_p(x, tab_fw2(s_tab, q, lane), 32, p);
}
__device__ __forceinline__ void inv_dev_rt(uint32_t (&x)[32], uint32_t* xbuf, const uint32_t* s_tab, int lane, int q, uint32_t p)
{
    pass_inv_p(x, c_iv1[q], 1, p);
    exchange_dev(x, xbuf, lane);
    pass_inv_p(x, tab_iv2(s_tab, q, lane), 32, p);
    const Tw* sc = tab_scale(s_tab, q, lane);
#pragma unroll
    for (int k = 0; k < 32; k++) {
        const Tw f = sc[k * 32];
        x[k] = shoup_mul_p(x[k], f.w, f.ws, p);
    }
}

// 32-bit shared-window address of a shared-memory pointer
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

}  // namespace b2
