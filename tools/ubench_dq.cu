// Realistic A/B of one in-register transform pass (four stages on register strides 8, 4, 2, 1 of two
// 16-coefficient polynomials, fifteen twiddles fetched from shared memory as 128-bit vectors -- spass_fwd<2>
// of ntt16s.cuh) with the Shoup quotient by IMAD.HI (today) against the quotient by one DFMA on the FP64 pipe:
//   q = low word of fma(B + v, ws, M - B ws),  B = 2^52 + 2^31 (v kept BIASED: its word with the top bit flipped IS
//   the low word of the double B + v), M = 1.5 * 2^84 (ulp 2^32: the mantissa's low word is round(v ws / 2^32)).
// 512 threads, 128 registers, one CTA per SM: the occupancy of k_blind_rotate_s.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/ubench_dq tools/ubench_dq.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITERS 512
constexpr int32_t P0 = 134215681;
struct alignas(16) V4 { int32_t x, y, z, w; };
struct Tws { int32_t w, ws; };
__device__ __forceinline__ V4 lds128(uint32_t a) { V4 r; asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a)); return r; }
__device__ __forceinline__ Tws lo(const V4& v) { return Tws{v.x, v.y}; }
__device__ __forceinline__ Tws hi(const V4& v) { return Tws{v.z, v.w}; }

// ---- today ----
struct TwA { int32_t w, ws; };
__device__ __forceinline__ TwA prepA(Tws t) { return TwA{t.w, t.ws}; }
__device__ __forceinline__ void bfly(int32_t& u, int32_t& v, const TwA t)
{
    const uint32_t q = (uint32_t)__mulhi(v, t.ws);
    const uint32_t s = (uint32_t)v * (uint32_t)t.w + (uint32_t)u;
    const uint32_t un = s - q * (uint32_t)P0;
    v = (int32_t)((uint32_t)u + (uint32_t)u - un);
    u = (int32_t)un;
}
// ---- quotient by DFMA (words biased by 2^31) ----
struct TwB { int32_t w; double wd, cw; };
__device__ __forceinline__ TwB prepB(Tws t)
{
    const double B = 4503599627370496.0 + 2147483648.0;          // 2^52 + 2^31
    const double M = 1.5 * 19342813113834066795298816.0;         // 1.5 * 2^84
    const double wd = __hiloint2double(0x43300000, t.ws) - B;    // table holds ws with the top bit flipped (even ws)
    return TwB{t.w, wd, fma(-B, wd, M)};
}
__device__ __forceinline__ void bfly(int32_t& u, int32_t& v, const TwB t)
{
    const double qd = fma(__hiloint2double(0x43300000, v), t.wd, t.cw);
    const uint32_t q = (uint32_t)__double2loint(qd);
    const uint32_t s = (uint32_t)v * (uint32_t)t.w + (uint32_t)u;
    const uint32_t un = s - q * (uint32_t)P0;
    v = (int32_t)((uint32_t)u + (uint32_t)u - un);
    u = (int32_t)un;
}
template <bool DQ, class TW> __device__ __forceinline__ void stage(int32_t (&x0)[16], int32_t (&x1)[16], const int tp, TW tw)
{
#pragma unroll
    for (int k = 0; k < 16; k++)
        if ((k & tp) == 0) {
            const Tws t = tw(k);
            if (DQ) { const TwB p = prepB(t); bfly(x0[k], x0[k + tp], p); bfly(x1[k], x1[k + tp], p); }
            else { const TwA p = prepA(t); bfly(x0[k], x0[k + tp], p); bfly(x1[k], x1[k + tp], p); }
        }
}
template <bool DQ> __device__ __forceinline__ void pass(int32_t (&x0)[16], int32_t (&x1)[16], const uint32_t a)
{
    const V4 c0 = lds128(a), c1 = lds128(a + 1024);
    stage<DQ>(x0, x1, 8, [&](int) { return lo(c0); });
    const V4 c2 = lds128(a + 2048), c3 = lds128(a + 3072);
    stage<DQ>(x0, x1, 4, [&](int k) { return (k >> 3) ? lo(c1) : hi(c0); });
    const V4 c4 = lds128(a + 4096), c5 = lds128(a + 5120), c6 = lds128(a + 6144), c7 = lds128(a + 7168);
    stage<DQ>(x0, x1, 2, [&](int k) { const int g = k >> 2; return g == 0 ? hi(c1) : g == 1 ? lo(c2) : g == 2 ? hi(c2) : lo(c3); });
    stage<DQ>(x0, x1, 1, [&](int k) { const int g = k >> 1; return g == 0 ? hi(c3) : g == 1 ? lo(c4) : g == 2 ? hi(c4) : g == 3 ? lo(c5) : g == 4 ? hi(c5) : g == 5 ? lo(c6) : g == 6 ? hi(c6) : lo(c7); });
}
template <bool DQ>
__global__ void __launch_bounds__(512, 1) k(uint32_t* out, uint32_t seed)
{
    __shared__ V4 tab[8][64];
    for (int i = threadIdx.x; i < 512; i += 512) tab[i >> 6][i & 63] = V4{(int32_t)(seed * i) % P0 & ~1, (int32_t)(seed * 7 * i) & ~1, (int32_t)(seed * 3 * i) % P0 & ~1, (int32_t)(seed * 5 * i) & ~1};
    __syncthreads();
    int32_t x0[16], x1[16];
#pragma unroll
    for (int i = 0; i < 16; i++) { x0[i] = (int32_t)(seed * (threadIdx.x + 1) + i); x1[i] = x0[i] ^ 0x5bd1e995; }
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(&tab[0][threadIdx.x & 63]);
    for (int it = 0; it < ITERS; it++) pass<DQ>(x0, x1, a);
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) acc ^= (uint32_t)x0[i] ^ (uint32_t)x1[i];
    if (acc == 0x12345678u) out[0] = acc;
}
template <bool DQ> void run(const char* name, uint32_t* d, int sms)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(a); k<DQ><<<sms, 512>>>(d, 12345u + r); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (r && ms < best) best = ms;
    }
    const double bf = (double)sms * 16 * ITERS * 64;                      // warp-level butterflies
    const double cyc = best * 1e-3 * 1.965e9 * sms * 4;
    printf("%-40s %8.3f ms   %.2f SMSP-cycles per warp-butterfly\n", name, best, cyc / bf);
}
int main()
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    uint32_t* d; cudaMalloc(&d, 16);
    run<false>("pass with IMAD.HI quotient (today)", d, sms);
    run<true>("pass with DFMA quotient", d, sms);
    return 0;
}
