// dev_common.cuh -- device-side plumbing shared by all kernels: named barriers, shared-window addresses.
#pragma once
#ifdef B2_HOST_TRACE
// Host build of the step schedule for tests/emul/emul_sched.cpp (a CPU race detector that runs the REAL
// brs_gate_loop source, one host thread per CUDA thread): no CUDA headers, the device qualifiers vanish, shared
// memory and every synchronisation go through the tracer the emulator defines.
#include <cstdint>
#include <cstddef>
#include <algorithm>
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __constant__
#define __restrict__
#define __launch_bounds__(...)
#define __cluster_dims__(...)
struct uint4 { uint32_t x, y, z, w; };
inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
struct TraceDim3 { unsigned x, y, z; };
extern thread_local TraceDim3 threadIdx, blockIdx, blockDim, gridDim;      // set by the emulator per host thread
template <class T> inline T __ldg(const T* p) { return *p; }
using std::min;
#include "ntt_core.cuh"
#include "tfhe_core.cuh"
namespace b2 {
uint32_t trace_smem_u32(const void* p);                   // byte offset inside the emulated shared memory (all CTAs)
uint32_t* trace_cta_smem();                               // my CTA's dynamic shared memory
void trace_bar(int id, int nthreads);                     // bar.sync id, nthreads (of my CTA)
void trace_warpsync();                                    // __syncwarp()
uint32_t trace_ld(uint32_t addr, int word);               // one 32-bit word of a 4 / 8 / 16-byte access at addr
void trace_st(uint32_t addr, int word, uint32_t v);
bool trace_sync_kept(int site);                           // false: the emulator drops synchronisation point `site`
void trace_mbar_init(uint32_t bar, uint32_t count);
void trace_mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes);
void trace_mbar_wait(uint32_t bar, uint32_t parity);
void trace_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar);
inline uint32_t smem_u32(const void* p) { return trace_smem_u32(p); }
inline void named_bar_sync(int id, int nthreads) { trace_bar(id, nthreads); }
inline void mbar_init(uint32_t bar, uint32_t count) { trace_mbar_init(bar, count); }
inline void mbar_fence_init() {}
inline void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) { trace_mbar_arrive_expect_tx(bar, bytes); }
inline void mbar_wait(uint32_t bar, uint32_t parity) { trace_mbar_wait(bar, parity); }
inline void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) { trace_bulk_g2s(dst, src, bytes, bar); }
}  // namespace b2
inline void __syncthreads() { b2::trace_bar(0, (int)blockDim.x); }
// dynamic shared memory of the kernel / plain-pointer accesses to it (traced on the host, plain on the device)
#define B2_DYN_SMEM(name) uint32_t* name = b2::trace_cta_smem()
#define B2_SMEM_RD(p) (b2::trace_ld(b2::trace_smem_u32(p), 0))
#define B2_SMEM_WR(p, v) (b2::trace_st(b2::trace_smem_u32(p), 0, (v)))
#else
#include <cuda_runtime.h>
#include "ntt_core.cuh"
#include "tfhe_core.cuh"
#define B2_DYN_SMEM(name) extern __shared__ __align__(16) uint32_t name[]
#define B2_SMEM_RD(p) (*(p))
#define B2_SMEM_WR(p, v) (*(p) = (v))

namespace b2 {

// 32-bit shared-window address of a shared-memory pointer
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- mbarrier + bulk-copy (TMA, 1-D) plumbing -----------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init()      // mbarrier.init visible to the async proxy / other CTAs
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// default (.acquire.cta) wait, as for TMA-written shared memory: the data is in shared memory when the
// transaction count completes; an .acquire.cluster wait would add an L1 invalidate (CCTL.IVALL)
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
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
// bytes (multiple of 16, 16-byte aligned on both sides) global -> my CTA's shared memory; completes on `bar`
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar) : "memory");
}

}  // namespace b2
#endif  // B2_HOST_TRACE
