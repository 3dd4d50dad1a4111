// Issue-port / pipe co-issue micro-benchmark for sm_100a: how many cycles does one IMAD.HI (or
// IMAD.WIDE, IMAD) plus k independent ALU instructions take per SM sub-partition?  If the ALU
// instructions were free next to the multiply, cycles would not grow with k.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_mix tools/ubench_mix.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITERS 1024
// MUL: 0 none, 1 IMAD, 2 IMAD.HI, 3 IMAD.WIDE (+xor to stay 32-bit), 4 FFMA ; NALU ALU ops of kind AK per multiply
// AK: 0 IADD3 (3 regs) 1 LOP3 2 SHF 3 IADD reg+imm (VIADD) 4 PRMT 5 ISETP+SEL 6 LEA 7 FADD 8 HADD2
template <int MUL, int NALU, int AK>
__global__ void __launch_bounds__(512) k(uint32_t* out, uint32_t seed)
{
    uint32_t x[8], y[8]; float f[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { x[i] = seed * (threadIdx.x + 1) + i; y[i] = x[i] ^ 0x5bd1e995u; f[i] = (float)x[i]; }
    const uint32_t w = seed | 1u, z = seed * 7u;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MUL == 1) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(w), "r"(z));
            if (MUL == 2) asm volatile("mul.hi.s32 %0, %0, %1;" : "+r"(x[i]) : "r"(w));
            if (MUL == 3) { uint32_t lo, hi; asm volatile("{ .reg .b64 t; mul.wide.s32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo), "=r"(hi) : "r"(x[i]), "r"(w)); x[i] = hi; y[i] ^= lo; }
            if (MUL == 4) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(1.0001f), "f"(0.5f));
#pragma unroll
            for (int a = 0; a < NALU; a++) {
                if (AK == 0) asm volatile("{ .reg .s32 t; add.s32 t, %0, %1; add.s32 %0, t, %2; }" : "+r"(y[i]) : "r"(w), "r"(z));   // fused by ptxas into IADD3
                if (AK == 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(y[i]) : "r"(w), "r"(z));
                if (AK == 2) asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(y[i]) : "r"(w));
                if (AK == 3) asm volatile("add.s32 %0, %0, 12345;" : "+r"(y[i]));
                if (AK == 4) asm volatile("prmt.b32 %0, %0, %1, 0x3120;" : "+r"(y[i]) : "r"(w));
                if (AK == 5) asm volatile("{ .reg .pred p; setp.gt.u32 p, %0, %1; selp.b32 %0, %2, %0, p; }" : "+r"(y[i]) : "r"(w), "r"(z));
                if (AK == 6) asm volatile("{ .reg .b32 t; shl.b32 t, %0, 5; add.s32 %0, t, %1; }" : "+r"(y[i]) : "r"(w));
                if (AK == 7) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(0.5f));
                if (AK == 8) asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(y[i]) : "r"(w));
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) acc ^= x[i] ^ y[i] ^ (uint32_t)f[i];
    if (acc == 0x12345678u) out[0] = acc;
}
template <int MUL, int NALU, int AK> void run(const char* name, uint32_t* d, int sms)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        cudaEventRecord(a); k<MUL, NALU, AK><<<sms * 4, 512>>>(d, 12345u + r); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (r && ms < best) best = ms;
    }
    // groups (one multiply + NALU alu) per SMSP per cycle: warps per SMSP = 4*512/32/4 = 16
    const double groups = (double)sms * 4 * 512 / 32 * ITERS * 8;        // warp-level groups
    const double cyc = best * 1e-3 * 1.965e9 * sms * 4;                  // SMSP-cycles
    printf("%-44s %8.3f ms   %.2f SMSP-cycles per warp-group\n", name, best, cyc / groups);
    fflush(stdout);
}
int main()
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    uint32_t* d; cudaMalloc(&d, 16);
    run<1, 0, 0>("IMAD", d, sms); run<2, 0, 0>("IMAD.HI", d, sms); run<3, 0, 0>("IMAD.WIDE", d, sms); run<4, 0, 0>("FFMA", d, sms);
    run<0, 1, 0>("IADD3", d, sms); run<0, 1, 1>("LOP3", d, sms); run<0, 1, 2>("SHF", d, sms); run<0, 1, 3>("add imm (VIADD)", d, sms);
    run<0, 1, 4>("PRMT", d, sms); run<0, 1, 5>("ISETP+SEL", d, sms); run<0, 1, 6>("shl+add (LEA)", d, sms); run<0, 1, 7>("FADD", d, sms); run<0, 1, 8>("HADD2", d, sms);
    run<1, 1, 0>("IMAD + 1 IADD3", d, sms); run<1, 2, 0>("IMAD + 2 IADD3", d, sms); run<1, 3, 0>("IMAD + 3 IADD3", d, sms);
    run<2, 1, 0>("IMAD.HI + 1 IADD3", d, sms); run<2, 2, 0>("IMAD.HI + 2 IADD3", d, sms); run<2, 3, 0>("IMAD.HI + 3 IADD3", d, sms); run<2, 4, 0>("IMAD.HI + 4 IADD3", d, sms);
    run<3, 1, 0>("IMAD.WIDE + 1 IADD3", d, sms); run<3, 2, 0>("IMAD.WIDE + 2 IADD3", d, sms); run<3, 3, 0>("IMAD.WIDE + 3 IADD3", d, sms);
    run<2, 2, 1>("IMAD.HI + 2 LOP3", d, sms); run<2, 2, 2>("IMAD.HI + 2 SHF", d, sms); run<2, 2, 3>("IMAD.HI + 2 VIADD", d, sms); run<2, 2, 7>("IMAD.HI + 2 FADD", d, sms); run<2, 2, 8>("IMAD.HI + 2 HADD2", d, sms);
    run<1, 1, 3>("IMAD + 1 VIADD", d, sms); run<1, 1, 7>("IMAD + 1 FADD", d, sms); run<1, 2, 7>("IMAD + 2 FADD", d, sms); run<1, 1, 8>("IMAD + 1 HADD2", d, sms);
    run<4, 1, 0>("FFMA + 1 IADD3", d, sms); run<4, 2, 0>("FFMA + 2 IADD3", d, sms);
    return 0;
}
