// dev_common.cuh -- device-side plumbing shared by all kernels: named barriers, shared-window addresses.
#pragma once
#include <cuda_runtime.h>
#include "ntt_core.cuh"
#include "tfhe_core.cuh"

namespace b2 {

// 32-bit shared-window address of a shared-memory pointer
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

}  // namespace b2
