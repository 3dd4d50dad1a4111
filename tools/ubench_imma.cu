// What the tensor cores of a B200 give a legacy int8 kernel: mma.sync.m16n8k32.s8 throughput (register operands,
// no memory traffic), the upper bound for an int8 limb formulation of the external product written without
// tcgen05.  DESIGN.md 2 compares it with the 42.7 G int8 operations per gate that formulation needs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_imma tools/ubench_imma.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITERS 4096
__global__ void __launch_bounds__(256) k(int* out, int seed)
{
    int a[4], b[2], c[8][4];
#pragma unroll
    for (int i = 0; i < 4; i++) a[i] = seed * (threadIdx.x + i + 1);
    b[0] = seed ^ threadIdx.x; b[1] = seed + threadIdx.x;
#pragma unroll
    for (int j = 0; j < 8; j++)
#pragma unroll
        for (int i = 0; i < 4; i++) c[j][i] = 0;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int j = 0; j < 8; j++)      // eight independent accumulators
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    int acc = 0;
#pragma unroll
    for (int j = 0; j < 8; j++)
#pragma unroll
        for (int i = 0; i < 4; i++) acc ^= c[j][i];
    if (acc == 0x12345678) out[0] = acc;
}
int main()
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int* d; cudaMalloc(&d, 16);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int bps : {1, 2, 4, 8}) {
        float best = 1e30f;
        for (int r = 0; r < 4; r++) {
            cudaEventRecord(a); k<<<sms * bps, 256>>>(d, 12345 + r); cudaEventRecord(b); cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b); if (r && ms < best) best = ms;
        }
        const double ops = (double)sms * bps * 8 /*warps*/ * ITERS * 8 * (16.0 * 8 * 32 * 2);
        printf("{\"probe\": \"imma_m16n8k32_s8\", \"ctas_per_sm\": %d, \"ms\": %.3f, \"int8_tops\": %.1f}\n", bps, best, ops / best * 1e-9);
    }
    return 0;
}
